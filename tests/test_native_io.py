"""CPU tests of the native host library: crc32c vectors, SSTable, TF1 bundle layout goldens
(SURVEY Appendix B), tfevents readable by tensorboard's own loader, idx reader vs python."""
import os
import struct

import numpy as np
import pytest

from distributedtensorflowexample_b200.utils import native_io as nio


def test_crc32c_known_vectors():
    assert nio.crc32c(b"") == 0
    assert nio.crc32c(b"123456789") == 0xE3069283
    assert nio.crc32c(b"\x00" * 32) == 0x8A9136AA           # RFC 3720 B.4
    assert nio.crc32c(b"\xff" * 32) == 0x62A8AB43
    assert nio.crc32c(bytes(range(32))) == 0x46DD794E
    m = nio.crc32c_masked(b"123456789")
    c = 0xE3069283
    assert m == (((c >> 15) | (c << 17)) + 0xA282EAD8) & 0xFFFFFFFF


def test_crc32c_matches_tensorboard_stub():
    tb = pytest.importorskip("tensorboard.compat.tensorflow_stub.pywrap_tensorflow")
    data = os.urandom(1000)
    assert nio.crc32c_masked(data) == tb.masked_crc32c(data)


@pytest.mark.parametrize("n,block", [(1, 4096), (40, 4096), (500, 256)])
def test_sstable_roundtrip(tmp_path, n, block):
    items = sorted((f"key/{i:05d}".encode(), os.urandom(i % 37)) for i in range(n))
    p = str(tmp_path / "t.sst")
    nio.sstable_write(p, items, block_size=block)
    assert nio.sstable_read(p) == items
    raw = open(p, "rb").read()
    assert raw[-8:] == struct.pack("<Q", 0xDB4775248B80FB57)
    # corrupting a data byte must be caught by the block crc
    bad = bytearray(raw); bad[5] ^= 0xFF
    open(p, "wb").write(bytes(bad))
    with pytest.raises(Exception):
        nio.sstable_read(p)


def test_sstable_rejects_unsorted(tmp_path):
    with pytest.raises(Exception):
        nio.sstable_write(str(tmp_path / "x"), [(b"b", b"1"), (b"a", b"2")])


def _reference_tensors(step=7):
    rng = np.random.default_rng(0)
    return {
        "global/dense/kernel": rng.normal(size=(784, 100)).astype(np.float32),
        "global/dense/bias": np.zeros(100, np.float32),
        "global/dense_1/kernel": rng.normal(size=(100, 10)).astype(np.float32),
        "global/dense_1/bias": np.arange(10, dtype=np.float32),
        "global/global_step": np.array(step, dtype=np.int32),
    }


def test_bundle_layout_golden(tmp_path):
    """Key order, offsets 0/400/314000/314040/318040, dtype codes, total 318,044 B (Appendix B)."""
    t = _reference_tensors()
    prefix = str(tmp_path / "model.ckpt-7")
    nio.bundle_write(prefix, t)
    assert os.path.getsize(prefix + ".data-00000-of-00001") == 318044
    ents = nio.bundle_entries(prefix)
    assert [e["name"] for e in ents] == ["global/dense/bias", "global/dense/kernel", "global/dense_1/bias",
                                         "global/dense_1/kernel", "global/global_step"]
    assert [e["offset"] for e in ents] == [0, 400, 314000, 314040, 318040]
    assert [e["size"] for e in ents] == [400, 313600, 40, 4000, 4]
    assert [e["dtype"] for e in ents] == [1, 1, 1, 1, 3]
    assert ents[1]["shape"] == [784, 100] and ents[4]["shape"] == []
    data = open(prefix + ".data-00000-of-00001", "rb").read()
    assert data[400:400 + 313600] == t["global/dense/kernel"].tobytes()
    assert struct.unpack("<i", data[318040:])[0] == 7
    for e in ents:
        assert e["crc32c"] == nio.crc32c_masked(data[e["offset"]:e["offset"] + e["size"]])
    # header entry: key "" -> num_shards=1, version.producer=1, little endian omitted (proto3 default)
    kv = nio.sstable_read(prefix + ".index")
    assert kv[0] == (b"", bytes([0x08, 0x01, 0x1A, 0x02, 0x08, 0x01]))
    # entry proto of the scalar: dtype=3, shape present-but-empty, offset, size, fixed32 crc
    gs = dict(kv)[b"global/global_step"]
    assert gs[:4] == bytes([0x08, 0x03, 0x12, 0x00])
    assert not os.path.exists(prefix + ".meta")


def test_bundle_roundtrip(tmp_path):
    t = _reference_tensors(123)
    prefix = str(tmp_path / "m")
    nio.bundle_write(prefix, t)
    back = nio.bundle_read(prefix)
    assert set(back) == set(t)
    for k in t:
        assert back[k].dtype == t[k].dtype and back[k].shape == t[k].shape
        np.testing.assert_array_equal(back[k], t[k])
    # corrupt tensor bytes -> crc error
    p = prefix + ".data-00000-of-00001"
    raw = bytearray(open(p, "rb").read()); raw[1000] ^= 1
    open(p, "wb").write(bytes(raw))
    with pytest.raises(Exception):
        nio.bundle_read(prefix)


def test_events_readable_by_tensorboard(tmp_path):
    loader = pytest.importorskip("tensorboard.backend.event_processing.event_file_loader")
    path = str(tmp_path / "events.out.tfevents.123.host")
    w = nio.EventsFile(path)
    w.write_version(100.0)
    w.write_session_log(100.5, 0, nio.SESSION_START)
    for s in range(5):
        w.write_scalars(101.0 + s, s, ["local/loss", "local/accuracy_1"], [9.75 - s, 0.1 * s])
    w.write_session_log(110.0, 4, nio.SESSION_CHECKPOINT, checkpoint_path="/x/model.ckpt-4")
    w.write_scalars(111.0, 4, ["global/global_step/sec"], [12345.5])
    w.flush()
    assert w.record_count == 9
    w.close()
    events = list(loader.LegacyEventFileLoader(path).Load())
    assert len(events) == 9
    assert events[0].file_version == "brain.Event:2" and events[0].wall_time == 100.0
    assert events[1].session_log.status == 1
    sc = events[2 + 3]
    assert sc.step == 3
    vals = {v.tag: v.simple_value for v in sc.summary.value}
    assert vals["local/loss"] == pytest.approx(6.75) and vals["local/accuracy_1"] == pytest.approx(0.3)
    assert events[7].session_log.status == 3 and events[7].session_log.checkpoint_path == "/x/model.ckpt-4"
    assert events[8].summary.value[0].tag == "global/global_step/sec"


def test_idx_native_matches_python(data_dir):
    from distributedtensorflowexample_b200.data import mnist
    import hashlib
    imgs = nio.read_idx(os.path.join(data_dir, mnist.TEST_IMAGES))
    labs = nio.read_idx(os.path.join(data_dir, mnist.TEST_LABELS))
    assert imgs.shape == (10000, 28, 28) and labs.shape == (10000,)
    np.testing.assert_array_equal(imgs, mnist.read_idx_python(os.path.join(data_dir, mnist.TEST_IMAGES)))
    np.testing.assert_array_equal(labs, mnist.read_idx_python(os.path.join(data_dir, mnist.TEST_LABELS)))
    assert np.bincount(labs).tolist() == [980, 1135, 1032, 1010, 982, 892, 958, 1028, 974, 1009]
    assert abs(float(imgs.mean()) - 33.79) < 0.01
    tl = nio.read_idx(os.path.join(data_dir, mnist.TRAIN_LABELS))
    assert tl.shape == (60000,)
    md5 = hashlib.md5(open(os.path.join(data_dir, mnist.TEST_LABELS), "rb").read()).hexdigest()
    assert md5.startswith("ec29112d")


def test_idx_rejects_garbage(tmp_path):
    import gzip
    p = tmp_path / "bad.gz"
    with gzip.open(p, "wb") as f:
        f.write(struct.pack(">II", 0x1234, 3) + b"abc")
    with pytest.raises(Exception):
        nio.read_idx(str(p))
    p2 = tmp_path / "short.gz"
    with gzip.open(p2, "wb") as f:
        f.write(struct.pack(">II", 0x801, 10) + b"abc")
    with pytest.raises(Exception):
        nio.read_idx(str(p2))


def test_events_batch_is_bytewise_the_per_step_writes(tmp_path):
    from distributedtensorflowexample_b200.utils.events import FileWriter, training_summary
    a = FileWriter(str(tmp_path / "a")); b = FileWriter(str(tmp_path / "b"))
    steps, losses, accs = [5, 6, 9], [1.5, 0.25, 3.0], [0.1, 0.5, 1.0]
    for s, l, ac in zip(steps, losses, accs):
        a.add_summary(training_summary(l, ac), s, wall_time=123.5)
    b.add_training_summaries(steps, losses, accs, wall_time=123.5)
    a.close(); b.close()
    da, db = open(a.path, "rb").read(), open(b.path, "rb").read()
    # the first record (file version) carries each writer's own creation time; everything after must match
    def tail(d):
        import struct
        n = struct.unpack("<Q", d[:8])[0]
        return d[8 + 4 + n + 4:]
    assert tail(da) == tail(db) and len(tail(da)) > 100


def test_events_array_writer_is_bytewise_the_per_step_writes(tmp_path):
    """The worker loop hands the summary thread numpy arrays (int32 steps, fp32 loss/accuracy of one launch)."""
    from distributedtensorflowexample_b200.utils.events import FileWriter, training_summary
    a = FileWriter(str(tmp_path / "a")); b = FileWriter(str(tmp_path / "b"))
    steps = np.array([5, 6, 9, 70000], np.int32)
    losses = np.array([1.5, 0.25, 3.0, 1e-3], np.float32); accs = np.array([0.1, 0.5, 1.0, 0.0], np.float32)
    for s, l, ac in zip(steps, losses, accs):
        a.add_summary(training_summary(float(l), float(ac)), int(s), wall_time=123.5)
    b.add_training_arrays(steps, losses, accs, wall_time=123.5)
    a.close(); b.close()
    import struct
    def tail(d):
        n = struct.unpack("<Q", d[:8])[0]
        return d[8 + 4 + n + 4:]
    da, db = open(a.path, "rb").read(), open(b.path, "rb").read()
    assert tail(da) == tail(db) and len(tail(da)) > 100


# ---- property tests (hypothesis): arbitrary contents must round-trip through the native formats -------------
from hypothesis import HealthCheck, given, settings, strategies as st

_SLOW_OK = settings(max_examples=40, deadline=None, suppress_health_check=[HealthCheck.function_scoped_fixture])


@_SLOW_OK
@given(items=st.dictionaries(st.binary(min_size=1, max_size=40), st.binary(max_size=300), max_size=60),
       block=st.sampled_from([64, 256, 4096]))
def test_sstable_roundtrip_property(tmp_path, items, block):
    pairs = sorted(items.items())
    p = str(tmp_path / "p.sst")
    nio.sstable_write(p, pairs, block_size=block)
    assert nio.sstable_read(p) == pairs


_NAME = st.text(alphabet="abcdefghijklmnopqrstuvwxyz_/0123456789", min_size=1, max_size=24)
_DTYPE = st.sampled_from([np.float32, np.float64, np.int32, np.int64])
_SHAPE = st.lists(st.integers(min_value=0, max_value=7), min_size=0, max_size=4)


@_SLOW_OK
@given(spec=st.dictionaries(_NAME, st.tuples(_DTYPE, _SHAPE), min_size=1, max_size=8), seed=st.integers(0, 2 ** 31 - 1))
def test_bundle_roundtrip_property(tmp_path, spec, seed):
    rng = np.random.default_rng(seed)
    tensors = {}
    for name, (dt, shape) in spec.items():
        a = rng.integers(-1000, 1000, size=shape) if np.issubdtype(dt, np.integer) else rng.normal(size=shape)
        tensors[name] = np.asarray(a, dtype=dt)
    prefix = str(tmp_path / "b")
    nio.bundle_write(prefix, tensors)
    back = nio.bundle_read(prefix)
    assert set(back) == set(tensors)
    for k, v in tensors.items():
        assert back[k].dtype == v.dtype and back[k].shape == v.shape
        np.testing.assert_array_equal(back[k], v)


@_SLOW_OK
@given(data=st.binary(max_size=5000))
def test_crc32c_incremental_equals_one_shot(data):
    cut = len(data) // 3
    assert nio.crc32c(data) == nio.crc32c_extend(nio.crc32c(data[:cut]), data[cut:])
