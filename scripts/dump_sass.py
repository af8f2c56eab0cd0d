"""Regenerate the trimmed SASS listings under profiles/sass/ from the built library (no GPU needed):

    python scripts/dump_sass.py

For each fused kernel named below: resource usage, the instruction histogram, and every instruction that proves
what the kernel fuses - tensor-core MMAs (UTCHMMA), TMA loads / reduce-stores (UTMALDG / UTMAREDG), TMEM loads
(LDTM), the push reductions (RED / REDG / REDUX at .SYS), system-scope loads and fences, cluster barriers - each with
a few lines of context, in program order.  The full listing is `cuobjdump -sass <lib>`.
"""
import os, re, subprocess, sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
LIB = os.path.join(ROOT, "distributedtensorflowexample_b200", "_lib", "libdtfe_cuda.so")
OUT = os.path.join(ROOT, "profiles", "sass")

# output name -> (substring of the demangled kernel name, what it is)
KERNELS = {
    "k_mlp2_step_8_async": ("k_mlp2_step<8, false>", "fused MNIST step, fp32 SIMT, lean instantiation: asynchronous training (pull = cp.async / ld.relaxed.sys from the shard, push = red.add.sys) and evaluation"),
    "k_mlp2_step_8_sync": ("k_mlp2_step<8, true>", "the same step with the synchronous exchange as phase D (owners' reduce-scatter, per-CTA-flag all-reduce, {value, step} pair transport)"),
    "k_mlp2_step_tc_10": ("k_mlp2_step_tc<10>", "fused MNIST step on tcgen05 kind::tf32 (8-CTA cluster, W1 pulled by TMA, X by TMA gather4)"),
    "k_gemm_tc2_bf16_fwd_bias_act": ("k_gemm_tc2<0, 0, 1, 1>", "2-CTA GEMM, forward X.W (B MN-major = shard weights by TMA) + bias + sigmoid epilogue"),
    "k_gemm_tc2_bf16_dw_push": ("k_gemm_tc2<0, 1, 1, 3>", "2-CTA GEMM, dW = X^T.dY with the push (red.add.v4.bf16x2) as its epilogue"),
    "k_gemm_tc2_bf16_dw_push_tma": ("k_gemm_tc2<0, 1, 1, 4>", "2-CTA GEMM, dW with the push as TMA reduce-adds (cp.reduce.async.bulk.tensor) for peer shards"),
    "k_push_sgd_f32": ("k_push_sgd_f32", "stand-alone push + apply (bandwidth sweep)"),
    "k_sync_rs_apply": ("k_sync_rs_apply", "stand-alone synchronous exchange: reduce-scatter fused with the apply"),
    "k_adam_apply": ("k_adam_apply", "owner-side Adam, atomic swap of the gradient mailbox"),
}
KEY = re.compile(r"\b(UTCHMMA|UTCQMMA|UTMALDG|UTMAREDG|UTMASTG|LDTM|UTCBAR|UCGABAR\w*|RED\b|REDG|ATOMG?|ATOM\b|MEMBAR|ERRBAR|"
                 r"LDG\.E\.\S*SYS|LDG\.E\.STRONG\.SYS|STG\.E\.\S*SYS|SYNCS|LDGSTS|UTMAPF|CCTL|FENCE)\b")


def main():
    os.makedirs(OUT, exist_ok=True)
    sass = subprocess.run(["cuobjdump", "-sass", LIB], capture_output=True, text=True, check=True).stdout
    res = subprocess.run(["cuobjdump", "-res-usage", LIB], capture_output=True, text=True).stdout
    blocks = re.split(r"\n\s*Function : ", sass)
    funcs = {}
    for b in blocks[1:]:
        name, _, body = b.partition("\n")
        dem = subprocess.run(["c++filt", name.strip()], capture_output=True, text=True).stdout.strip()
        funcs[dem] = (name.strip(), body)
    usage = {}
    for m in re.finditer(r"Function (\S+):\s*\n\s*(REG:[^\n]*)", res):
        usage[m.group(1)] = m.group(2).strip()
    for out_name, (needle, what) in KERNELS.items():
        match = [d for d in funcs if needle in d]
        if not match:
            print(f"{out_name}: no kernel matching {needle!r}", file=sys.stderr)
            continue
        dem = match[0]
        mangled, body = funcs[dem]
        lines = [l for l in body.splitlines() if re.match(r"\s+/\*[0-9a-f]{4,}\*/", l)]
        ins = [re.sub(r"\s+/\*[0-9a-f]{16}\*/\s*$", "", l).rstrip() for l in lines]
        hist = {}
        for l in ins:
            m = re.match(r"\s+/\*[0-9a-f]+\*/\s+(?:@!?U?P\w+\s+)?([A-Z0-9_]+(?:\.[A-Z0-9_]+)*)", l)
            if m:
                op = m.group(1)
                head = ".".join(op.split(".")[:3]) if op.startswith(("UTC", "UTMA", "LDTM", "RED", "ATOM", "LDG", "STG", "MEMBAR")) else op.split(".")[0]
                hist[head] = hist.get(head, 0) + 1
        keep = set()
        for i, l in enumerate(ins):
            if KEY.search(l):
                keep.update(range(max(0, i - 2), min(len(ins), i + 3)))
        with open(os.path.join(OUT, out_name + ".sass.txt"), "w") as f:
            f.write(f"{dem}\n{what}\nsm_100a, {len(ins)} instructions; {usage.get(mangled, '')}\n\n")
            f.write("instruction histogram (tensor-core / TMA / TMEM / reduction / system-scope opcodes in full, others by base name):\n")
            for op, n in sorted(hist.items(), key=lambda kv: -kv[1])[:48]:
                f.write(f"  {n:6d} {op}\n")
            f.write("\nexcerpts around the instructions that carry the fusion, in program order ('...' = lines left out):\n")
            prev = -2
            for i in sorted(keep):
                if i != prev + 1:
                    f.write("        ...\n")
                f.write(ins[i] + "\n")
                prev = i
        print(f"{out_name}: {len(ins)} instructions, {len(keep)} kept -> profiles/sass/{out_name}.sass.txt")


if __name__ == "__main__":
    main()
