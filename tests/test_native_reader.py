"""Host-side input runtime (csrc/host, libkfb200_host.so): CRC32C, threaded TFRecord streaming,
tf.Example scanner - checked against the pure-Python codec of data/tfrecord.py."""
import os

import numpy as np
import pytest

from kf_benchmarks_b200 import build as build_lib
from kf_benchmarks_b200.data import native_reader, tfrecord


@pytest.fixture(scope="module", autouse=True)
def _built():
    build_lib.build_host()
    assert native_reader.available()


def test_crc32c_matches_python_and_known_vector():
    assert native_reader.crc32c(b"123456789") == 0xE3069283          # standard check value
    rng = np.random.RandomState(0)
    for n in (0, 1, 7, 8, 9, 63, 1000, 65537):
        data = rng.bytes(n)
        assert native_reader.crc32c(data) == tfrecord.crc32c(data)
        assert native_reader.masked_crc32c(data) == tfrecord.masked_crc(data)


def _write(tmp_path, nfiles, per_file):
    rng = np.random.RandomState(1)
    paths, all_recs = [], []
    for f in range(nfiles):
        recs = [bytes([f]) + rng.bytes(int(rng.randint(0, 3000))) for _ in range(per_file)]
        path = str(tmp_path / ("part-%05d" % f))
        tfrecord.write_records(path, recs)
        paths.append(path)
        all_recs.append(recs)
    return paths, all_recs


def test_single_thread_preserves_file_order(tmp_path):
    paths, recs = _write(tmp_path, 3, 20)
    got = list(native_reader.RecordReader(paths, num_threads=1))
    assert got == [r for f in recs for r in f]
    assert got == [r for p in paths for r in tfrecord.read_records(p)]


def test_threads_interleave_but_keep_per_file_order(tmp_path):
    paths, recs = _write(tmp_path, 5, 40)
    got = list(native_reader.RecordReader(paths, num_threads=3, capacity=8))
    assert sorted(got) == sorted(r for f in recs for r in f)
    for f in range(5):                                   # order within one file is preserved
        assert [r for r in got if r[0] == f] == recs[f]


def test_repeat_and_close_do_not_hang(tmp_path):
    paths, recs = _write(tmp_path, 2, 5)
    rd = native_reader.RecordReader(paths, num_threads=2, capacity=4, repeat=True,
                                    shuffle_files=True, seed=3)
    got = [next(rd) for _ in range(37)]                  # more than one pass
    assert len(got) == 37 and set(got) <= set(r for f in recs for r in f)
    rd.close()                                           # readers blocked on a full queue exit
    with pytest.raises(StopIteration):
        next(rd)


def test_corruption_is_reported(tmp_path):
    paths, recs = _write(tmp_path, 1, 6)
    raw = bytearray(open(paths[0], "rb").read())
    raw[len(raw) // 2] ^= 0x40                           # flip one bit in the middle
    open(paths[0], "wb").write(bytes(raw))
    rd = native_reader.RecordReader(paths, num_threads=1)
    with pytest.raises(IOError) as e:
        list(rd)
    assert "corrupt" in str(e.value)
    ok = list(native_reader.RecordReader(paths, num_threads=1, check_crc=False))
    assert len(ok) == 6                                  # without validation the bytes pass
    with pytest.raises(IOError):
        list(native_reader.RecordReader([str(tmp_path / "missing")]))


def test_example_scanner_matches_python_parser():
    feats = {"image/encoded": [b"\xff\xd8jpegbytes\x00\x01"], "image/class/label": [417],
             "image/object/bbox/xmin": [0.1, 0.25, 0.5], "image/class/text": [b"dog"],
             "negatives": [-3, 7, -(1 << 40)], "image/source_id": [b"12345"]}
    rec = tfrecord.encode_example(feats)
    assert native_reader.example_bytes(rec, "image/encoded") == feats["image/encoded"][0]
    assert native_reader.example_bytes(rec, "image/source_id") == b"12345"
    assert native_reader.example_numbers(rec, "image/class/label") == [417.0]
    assert native_reader.example_numbers(rec, "negatives") == [-3.0, 7.0, float(-(1 << 40))]
    assert np.allclose(native_reader.example_numbers(rec, "image/object/bbox/xmin"),
                       [0.1, 0.25, 0.5])
    assert native_reader.example_numbers(rec, "absent") is None
    assert native_reader.example_bytes(rec, "absent") is None
    assert native_reader.example_bytes(rec, "image/class/label") is None      # not a bytes list
    assert native_reader.example_numbers(rec, "negatives", cap=1) == [-3.0, 7.0,
                                                                      float(-(1 << 40))]
    py = tfrecord.parse_example(rec)
    for k, v in feats.items():
        assert py[k] == pytest.approx(v) if isinstance(v[0], float) else py[k] == v


# ---- sanitizer builds of the host runtime (TSAN build of the C++ runtime) -------
def _sanitizer_run(tmp_path, name, flags, env_extra):
    import glob
    import shutil
    import subprocess
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    srcs = sorted(glob.glob(os.path.join(root, "csrc", "host", "*.cpp")))
    srcs.append(os.path.join(root, "csrc", "host", "stress", "reader_stress.cpp"))
    exe = str(tmp_path / name)
    # $CXX may point at a toolchain that ships without the sanitizer runtimes: try the system
    # compilers as well before giving up
    res = None
    for cxx in dict.fromkeys([os.environ.get("CXX", "g++"), "/usr/bin/g++", "g++", "clang++"]):
        if shutil.which(cxx) is None:
            continue
        res = subprocess.run([cxx, "-g", "-O1", "-std=c++17", "-pthread", "-msse4.2"] + flags
                             + srcs + ["-o", exe], capture_output=True, text=True)
        if res.returncode == 0 or "cannot find" not in res.stderr:
            break
    if res is None:
        pytest.skip("no C++ compiler")
    if res.returncode != 0 and ("cannot find" in res.stderr or "unrecognized" in res.stderr):
        pytest.skip("sanitizer runtime not installed: " + res.stderr.strip()[-300:])
    assert res.returncode == 0, res.stderr
    scratch = tmp_path / "data"
    scratch.mkdir()
    env = dict(os.environ, **env_extra)
    run = subprocess.run([exe, str(scratch), "6"], capture_output=True, text=True, env=env,
                         timeout=600)
    report = run.stdout + run.stderr
    assert run.returncode == 0, report
    assert "reader_stress ok" in run.stdout
    assert "ThreadSanitizer" not in report and "AddressSanitizer" not in report \
        and "runtime error" not in report, report


def test_host_runtime_is_clean_under_thread_sanitizer(tmp_path):
    _sanitizer_run(tmp_path, "rs_tsan", ["-fsanitize=thread"],
                   {"TSAN_OPTIONS": "halt_on_error=1 exitcode=66"})


def test_host_runtime_is_clean_under_address_and_ub_sanitizers(tmp_path):
    _sanitizer_run(tmp_path, "rs_asan",
                   ["-fsanitize=address,undefined", "-fno-sanitize-recover=undefined"],
                   {"ASAN_OPTIONS": "detect_leaks=1"})


def test_deterministic_interleave_is_round_robin_over_files(tmp_path):
    """deterministic=True: one record from each reader in turn (tf.data interleave, block
    length 1); exhausted files drop out of the rotation; identical on every run."""
    rng = np.random.RandomState(5)
    paths, recs = [], []
    for f, n in enumerate([7, 3, 5]):
        r = [bytes([f, i]) + rng.bytes(int(rng.randint(0, 500))) for i in range(n)]
        path = str(tmp_path / ("part-%d" % f))
        tfrecord.write_records(path, r)
        paths.append(path)
        recs.append(r)
    want = []
    for i in range(7):
        for f in range(3):
            if i < len(recs[f]):
                want.append(recs[f][i])
    for _ in range(5):
        got = list(native_reader.RecordReader(paths, num_threads=3, capacity=6, deterministic=True))
        assert got == want
    # two readers over three files: reader 0 owns files 0 and 2 (one after the other), reader 1
    # owns file 1
    got = list(native_reader.RecordReader(paths, num_threads=2, capacity=4, deterministic=True))
    seq0, seq1 = recs[0] + recs[2], recs[1]
    want2 = []
    for i in range(max(len(seq0), len(seq1))):
        if i < len(seq0):
            want2.append(seq0[i])
        if i < len(seq1):
            want2.append(seq1[i])
    assert got == want2
    # repeat + early close does not hang
    rd = native_reader.RecordReader(paths, num_threads=3, capacity=3, repeat=True,
                                    deterministic=True)
    first = [next(rd) for _ in range(40)]
    rd.close()
    # every file restarts when it ends, so rotation r takes record r % len(file) of each file
    assert first == [recs[k % 3][(k // 3) % len(recs[k % 3])] for k in range(40)]
