"""Flag system with the reference's surface (reference main.py:26-39, worker.py:5).

``tf.app.flags`` semantics that the reference depends on and that are kept here:

* one flat, process-global namespace (``FLAGS``), defaults in code;
* lazy parsing: the command line is parsed on the first attribute access, because the
  reference calls ``main()`` directly and never ``tf.app.run()`` (main.py:79-80);
* both ``--key value`` and ``--key=value`` spellings (run_single_gpu.sh:17 uses both);
* unknown flags are tolerated and kept in ``FLAGS.unparsed``.

New flags (num_ps, sync, optimizer, dtype, data, hidden, ...) are additive; the reference's
eight keep their names and defaults.
"""
from __future__ import annotations

import os
import sys
from typing import Any, Callable, Dict, List, Optional


class _Flag:
    __slots__ = ("name", "default", "help", "parser", "kind")

    def __init__(self, name: str, default: Any, help: str, parser: Callable[[str], Any], kind: str):
        self.name, self.default, self.help, self.parser, self.kind = name, default, help, parser, kind


def _parse_bool(text: str) -> bool:
    t = text.strip().lower()
    if t in ("1", "true", "t", "yes", "y", "on"):
        return True
    if t in ("0", "false", "f", "no", "n", "off"):
        return False
    raise ValueError(f"not a boolean: {text!r}")


class FlagValues:
    """Lazy-parsing flag container (the object behind ``flags.FLAGS``)."""

    def __init__(self) -> None:
        object.__setattr__(self, "_defs", {})      # name -> _Flag
        object.__setattr__(self, "_values", {})    # name -> parsed/overridden value
        object.__setattr__(self, "_parsed", False)
        object.__setattr__(self, "unparsed", [])

    # -- definition -------------------------------------------------------------------
    def _define(self, flag: _Flag) -> None:
        self._defs[flag.name] = flag
        if self._parsed:  # defined after first access: re-scan the unparsed tail
            object.__setattr__(self, "_parsed", False)

    # -- parsing ----------------------------------------------------------------------
    def parse(self, argv: Optional[List[str]] = None) -> List[str]:
        """Parse ``argv`` (default ``sys.argv[1:]``); returns the tokens that were not flags."""
        args = list(sys.argv[1:] if argv is None else argv)
        rest: List[str] = []
        i = 0
        while i < len(args):
            tok = args[i]
            i += 1
            if tok == "--":
                rest.extend(args[i:])
                break
            if not tok.startswith("-") or tok in ("-", "--"):
                rest.append(tok)
                continue
            if tok in ("-h", "--help", "--helpfull", "--helpshort"):
                # tf.app.flags is argparse underneath: asking for help prints the flags and exits 0
                # (without this, `python main.py --help` would start the default role - a ps that never returns)
                print(f"usage: {os.path.basename(sys.argv[0] or 'main.py')} [flags]\n\nflags:\n{self.usage()}")
                raise SystemExit(0)
            body = tok.lstrip("-")
            if "=" in body:
                name, text = body.split("=", 1)
                has_value = True
            else:
                name, text, has_value = body, "", False
            name = name.replace("-", "_")
            negated = False
            if name not in self._defs and name.startswith("no") and name[2:] in self._defs \
                    and self._defs[name[2:]].kind == "bool":
                name, negated = name[2:], True
            flag = self._defs.get(name)
            if flag is None:
                rest.append(tok)   # unknown flag: tolerated, like tf.app.flags without tf.app.run()
                continue
            if flag.kind == "bool":
                if has_value:
                    self._values[name] = flag.parser(text)
                elif i < len(args) and not args[i].startswith("-") and _looks_bool(args[i]) and not negated:
                    self._values[name] = flag.parser(args[i])
                    i += 1
                else:
                    self._values[name] = not negated
                continue
            if not has_value:
                if i >= len(args):
                    raise ValueError(f"flag --{name} needs a value")
                text = args[i]
                i += 1
            try:
                self._values[name] = flag.parser(text)
            except ValueError as e:
                raise ValueError(f"flag --{name}={text!r}: {e}") from None
        object.__setattr__(self, "_parsed", True)
        object.__setattr__(self, "unparsed", rest)
        return rest

    def _ensure(self) -> None:
        if not self._parsed:
            self.parse()

    # -- access -----------------------------------------------------------------------
    def __getattr__(self, name: str) -> Any:
        defs = object.__getattribute__(self, "_defs")
        if name not in defs:
            raise AttributeError(f"unknown flag {name!r}")
        self._ensure()
        vals = object.__getattribute__(self, "_values")
        return vals[name] if name in vals else defs[name].default

    def __setattr__(self, name: str, value: Any) -> None:
        if name not in self._defs:
            raise AttributeError(f"unknown flag {name!r}")
        self._values[name] = value

    def __contains__(self, name: str) -> bool:
        return name in self._defs

    def flag_values_dict(self) -> Dict[str, Any]:
        self._ensure()
        return {n: (self._values[n] if n in self._values else f.default) for n, f in self._defs.items()}

    def reset(self) -> None:
        """Forget parsed/overridden values (tests)."""
        self._values.clear()
        object.__setattr__(self, "_parsed", False)
        object.__setattr__(self, "unparsed", [])

    def usage(self) -> str:
        lines = []
        for n, f in sorted(self._defs.items()):
            lines.append(f"  --{n} ({f.kind}, default {f.default!r}): {f.help}")
        return "\n".join(lines)


def _looks_bool(text: str) -> bool:
    try:
        _parse_bool(text)
        return True
    except ValueError:
        return False


FLAGS = FlagValues()


def DEFINE_string(name: str, default: Optional[str], help: str, flag_values: FlagValues = FLAGS) -> None:
    flag_values._define(_Flag(name, default, help, str, "string"))


def DEFINE_integer(name: str, default: Optional[int], help: str, flag_values: FlagValues = FLAGS) -> None:
    flag_values._define(_Flag(name, default, help, lambda s: int(s, 0) if not _is_float_int(s) else int(float(s)), "int"))


def DEFINE_float(name: str, default: Optional[float], help: str, flag_values: FlagValues = FLAGS) -> None:
    flag_values._define(_Flag(name, default, help, float, "float"))


def DEFINE_boolean(name: str, default: Optional[bool], help: str, flag_values: FlagValues = FLAGS) -> None:
    flag_values._define(_Flag(name, default, help, _parse_bool, "bool"))


DEFINE_bool = DEFINE_boolean


def _is_float_int(s: str) -> bool:
    # accept "1e7" style for training_steps
    return ("e" in s.lower() and not s.lower().startswith("0x"))


def define_reference_flags(flag_values: FlagValues = FLAGS) -> FlagValues:
    """The eight flags of the reference (main.py:28-37), same names, same defaults."""
    DEFINE_string("job_name", "ps", "Either 'ps' or 'worker'", flag_values)
    DEFINE_integer("task_index", 0, "Index of task within the job", flag_values)
    DEFINE_integer("batch_size", 100, "Batch size", flag_values)
    DEFINE_float("learning_rate", 0.001, "Learning rate", flag_values)
    DEFINE_integer("training_steps", 10 ** 7, "Training steps (1 step = 1 batch update)", flag_values)
    DEFINE_string("logdir", "./tmp/mnist/1", "Log directory", flag_values)
    DEFINE_integer("num_workers", 2, "Number of workers", flag_values)
    DEFINE_integer("num_gpus", 1, "Number of gpus, less than or equal to num_workers", flag_values)
    return flag_values


def define_framework_flags(flag_values: FlagValues = FLAGS) -> FlagValues:
    """Additive flags of this framework."""
    DEFINE_integer("num_ps", 1, "Number of parameter-server tasks (shards)", flag_values)
    DEFINE_boolean("sync", False, "Synchronous SGD (reduce-scatter fused with the update) instead of async", flag_values)
    DEFINE_string("optimizer", "sgd", "sgd | adam", flag_values)
    DEFINE_string("sync_mode", "allreduce_ll", "--sync with the fused step kernel: allreduce_ll (gradient elements travel as "
                  "{value, step} pairs into every rank's receive buffer) | allreduce (per-CTA flags + peer loads)", flag_values)
    DEFINE_string("device", "auto", "auto | cuda | cpu (cpu = gloo plumbing config)", flag_values)
    DEFINE_string("dtype", "float32", "float32 | bfloat16 parameter/compute dtype of the GPU path", flag_values)
    DEFINE_boolean("tf32", False, "Let the fp32 layer-wise engine use tcgen05 kind::tf32 GEMMs (TMA pull from the fp32 "
                   "shards) where operand pitches allow; default is exact fp32", flag_values)
    DEFINE_string("engine", "auto", "GPU step engine: auto | fused (fp32 SIMT persistent kernel) | fused_tc (the same step on tcgen05 kind::tf32) | layerwise | tc", flag_values)
    DEFINE_string("data", "mnist", "mnist | synthetic", flag_values)
    DEFINE_string("data_dir", "MNIST_data", "Directory with the MNIST idx.gz files", flag_values)
    DEFINE_string("hidden", "100", "Comma-separated hidden layer widths (reference: 100)", flag_values)
    DEFINE_integer("seed", -1, "RNG seed (-1 = unseeded like the reference)", flag_values)
    DEFINE_float("save_model_secs", 30, "Chief checkpoint period in seconds (reference: 30)", flag_values)
    DEFINE_float("save_summaries_secs", 30, "Chief global_step/sec summary period (reference: 30)", flag_values)
    DEFINE_integer("log_every", 100, "Print step/cost/speed when global step %% this == 0 (reference: 100)", flag_values)
    DEFINE_integer("eval_every", 10000, "Test-set accuracy when global step %% this == 0 (reference: 10000)", flag_values)
    DEFINE_integer("steps_per_launch", 0, "Training steps fused into one persistent kernel launch; 0 = adaptive (up to "
                   "256, shrinking towards evaluation steps and training_steps), 1 = one launch per step like one "
                   "sess.run per step", flag_values)
    DEFINE_string("ps_strategy", "round_robin", "round_robin | greedy_bytes variable placement", flag_values)
    DEFINE_boolean("log_device_placement", False, "Print where every variable was placed (TF1 ConfigProto option; "
                   "the reference sets it to False)", flag_values)
    DEFINE_integer("base_port", 12222, "First port of the cluster spec (reference: 12222)", flag_values)
    DEFINE_string("gpu_ids", "", "Comma-separated physical GPU ids (the launchers' GPU_ID array)", flag_values)
    DEFINE_boolean("unlocked_apply", False, "Apply pushes as unlocked read-modify-write like TF1 use_locking=False "
                   "(default: lossless red.add)", flag_values)
    DEFINE_string("fault", "", "Test-only fault injection, e.g. 'kill_worker:1@50' or 'delay_push:0.01'", flag_values)
    DEFINE_string("trace_file", "", "Write a Chrome-trace JSON of this task's host spans and CUDA-event device spans "
                  "(job and task index are appended to the name); also emits NVTX ranges", flag_values)
    DEFINE_float("sync_timeout_secs", 30, "Synchronous mode: give up on a round (and stop) after waiting this long "
                 "for the other workers", flag_values)
    DEFINE_float("recovery_wait_secs", 30, "Non-chief poll period while waiting for the chief's init (TF1: 30)", flag_values)
    return flag_values
