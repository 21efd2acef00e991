"""CPU: host logic of the drop-in surface (datasets, config checks, chunk planner, sharding,
C-ABI exports) and the world_size-2 gloo path."""
import ctypes
import json
import os
import re
import subprocess
import sys

import numpy as np
import pytest

import oracle_np as A
from nanopore_rnn_b200 import _lib, dataset, engine, run_nanotensor, shard, synthetic as syn
from nanopore_rnn_b200.error import KernelLibraryMissing, Usage

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_cabi_library_exports_every_declared_symbol():
    hdr = open(os.path.join(ROOT, "include", "nanotensor_b200.h")).read()
    declared = set(re.findall(r"NTB_API[^;(]*?\b(ntb_\w+)\s*\(", hdr))
    assert len(declared) >= 20
    assert declared == set(_lib.SIGNATURES), declared ^ set(_lib.SIGNATURES)
    lib = ctypes.CDLL(_lib.LIB_PATH)
    for name in declared:
        assert hasattr(lib, name), name
    assert b"sm_100a" in _lib.load().ntb_version()


def test_no_cpu_fallback_without_device():
    import torch
    if torch.cuda.is_available():
        pytest.skip("device present")
    lib = _lib.load()
    assert lib.ntb_device_ok() == 0
    rc = lib.ntb_decode(ctypes.c_void_p(8), ctypes.c_void_p(8), 1, 1, 2, 0, ctypes.c_void_p(8), None, None)
    assert rc == -2 and b"CUDA" in lib.ntb_last_error()
    net = syn.NETWORKS["event_fixture"]
    w = syn.random_weights(net, 12, 1025, 0, graph="CrossEntropy")
    with pytest.raises(_lib.NtbError):
        engine.Model(net, w, "CrossEntropy", 30, 12, 1025)


def test_missing_library_fails_loudly(monkeypatch):
    monkeypatch.setattr(_lib, "_lib", None)
    monkeypatch.setattr(_lib, "LIB_PATH", "/nonexistent/libnanotensor_b200.so")
    with pytest.raises(KernelLibraryMissing):
        _lib.load()


def test_plan_chunks_whole_groups():
    goff = engine.group_offsets([1000] * 30 + [7], 10)
    chunks = engine.plan_chunks(goff, 9472)
    assert chunks[0][0] == 0 and chunks[-1][1] == len(goff) - 1
    for (a, b), (c, d) in zip(chunks, chunks[1:]):
        assert b == c
    sizes = [int(goff[b] - goff[a]) for a, b in chunks]
    assert max(sizes) <= 9472 and sum(sizes) == 30007
    assert sizes[0] == 9470                       # largest multiple of the group size under 2 waves
    # aligned cuts (Model.chunk_align(): the fused BN statistics stay bit-identical under re-chunking)
    chunks = engine.plan_chunks(goff, 9472, align=16)
    sizes = [int(goff[b] - goff[a]) for a, b in chunks]
    assert sum(sizes) == 30007 and max(sizes) <= 9472
    assert all(int(goff[b]) % 16 == 0 for a, b in chunks[:-1])
    assert sizes[0] == 9440                       # largest multiple of lcm(10, 16) = 80 windows under the limit
    # no aligned boundary available: falls back to plain group boundaries
    odd = engine.group_offsets([25] * 8, 20)
    chunks = engine.plan_chunks(odd, 64, align=16)
    assert [int(odd[b] - odd[a]) for a, b in chunks] == [50, 50, 50, 50]
    with pytest.raises(ValueError):
        engine.plan_chunks(np.array([0, 100, 200]), 64)


def test_signal_dataset_windows_and_job(tmp_path):
    rng = np.random.default_rng(0)
    files = []
    for i, n in enumerate([950, 123]):
        p = tmp_path / ("read%d.signal" % i)
        p.write_text(" ".join(str(int(v)) for v in rng.integers(300, 700, n)))
        files.append(str(p))
    ds = dataset.FullSignalSequence(files, mode=2, batch_size=4, seq_len=100, step=100, verbose=False,
                                    inference_output_dir=str(tmp_path))
    x, lens, goff, read_off = ds.inference_job()
    assert x.shape == (12, 100) and lens.tolist() == [100] * 9 + [50] + [100, 23]
    assert goff.tolist() == [0, 4, 8, 10, 12] and read_off.tolist() == [0, 10, 12]
    raw = np.array(open(files[0]).read().split(), dtype=np.int64)
    xo, lo = A.windows_of_read(raw, 100, step=100)
    np.testing.assert_array_equal(x[:10], xo)
    # iterator = the reference's sess.run granularity
    ds.file_path = files[1]
    ds.iterator.initializer()
    bx, bl = ds.iterator.get_next()
    assert bx.shape == (2, 100) and bl.tolist() == [100, 23]
    with pytest.raises(dataset.OutOfRangeError):
        ds.iterator.get_next()
    out = ds.process_output([np.array([[0, 1, 2, 3, -1], [1, 2, 3, 0, 0]])], files[0])
    assert open(out).read() == ">read0\nACGTAA\n"
    # default step is 300 (dataset.py:309)
    assert dataset.FullSignalSequence(files, mode=2, verbose=False).step == 300
    with pytest.raises(AssertionError):
        dataset.FullSignalSequence(files, mode=2, batch_size="4", verbose=False)


def test_event_dataset(tmp_path):
    rng = np.random.default_rng(1)
    arr = np.empty((65, 2), dtype=object)
    for i in range(65):
        arr[i, 0] = rng.standard_normal(12)
        one = np.zeros(1025)
        one[rng.integers(0, 1025)] = 1
        arr[i, 1] = one
    np.save(str(tmp_path / "a.npy"), arr, allow_pickle=True)
    (tmp_path / "bad.npy").write_bytes(b"nope")
    ds = dataset.NumpyEventData([str(tmp_path / "a.npy"), str(tmp_path / "bad.npy")], mode=1, batch_size=2,
                                seq_len=30, verbose=False)
    assert ds.len_x == 12 and ds.len_y == 1025 and len(ds.bad_files) == 1
    x, lens, goff, y = ds.test_job()
    assert x.shape == (2, 30, 12) and y.shape == (2, 30) and goff.tolist() == [0, 2]
    xo, yo, _ = A.event_windows(arr, 30)
    np.testing.assert_allclose(x, xo.astype(np.float32))
    np.testing.assert_array_equal(y, yo.argmax(2))


def test_config_checks(tmp_path):
    cfg = json.load(open(os.path.join(ROOT, "tests", "golden", "configs", "signal_CRNN.config.json")))
    with pytest.raises(Usage):                                  # shipped config says "train": true
        run_nanotensor.CommandLine.check_args(cfg)
    cfg.update(train=False, inference=True, test=False)
    with pytest.raises(Usage):                                  # its directories do not exist here
        run_nanotensor.CommandLine.check_args(cfg)
    cfg["CreateDataset"]["inference_dir"] = str(tmp_path)
    cfg["CreateDataset"]["dataset_args"]["inference_output_dir"] = str(tmp_path)
    cfg["use_checkpoint"] = True
    cfg["trained_model"] = os.path.join(ROOT, "tests", "golden")
    args = run_nanotensor.CommandLine.check_args(cfg)
    assert args.CreateDataset.dataset == "FullSignalSequence" and args.BuildGraph.graph == "CtcLoss"
    cfg["train"] = True
    with pytest.raises(Usage):
        run_nanotensor.CommandLine.check_args(cfg)


def test_lpt_sharding_is_balanced_and_complete():
    lens = syn.lognormal_read_lengths(4096, seed=20)
    counts = (lens + 99) // 100
    for ws in (1, 2, 4, 8):
        a = shard.assign_reads(counts, ws)
        assert sorted(i for part in a for i in part) == list(range(len(counts)))
        loads = shard.shard_loads(counts, a)
        assert max(loads) - min(loads) <= counts.max()
    assert shard.assign_reads([5, 3, 9, 1, 1, 7], 2) == [[1, 2, 3], [0, 4, 5]]


def test_world_size_2_gloo_shard_and_gather(tmp_path):
    script = tmp_path / "w.py"
    script.write_text("""
import os, sys
sys.path.insert(0, %r)
import torch.distributed as dist
import nanopore_rnn_b200
from nanopore_rnn_b200 import shard
dist.init_process_group("gloo")
r, w = dist.get_rank(), dist.get_world_size()
counts = [5, 3, 9, 1, 1, 7, 2]
mine = shard.assign_reads(counts, w)[r]
local = {i: "read%%d@%%d" %% (i, r) for i in mine}
merged = shard.gather_labels(local, r, w)
assert sorted(merged) == list(range(len(counts))), merged
assert all(merged[i].endswith("@%%d" %% r) for i in mine)
dist.barrier()
if r == 0:
    print("OK", len(merged))
dist.destroy_process_group()
""" % ROOT)
    out = subprocess.run([sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node=2",
                          "--master-addr", "127.0.0.1", "--master-port", "29617", str(script)],
                         capture_output=True, text=True, timeout=240)
    assert out.returncode == 0, out.stderr[-2000:]
    assert "OK 7" in out.stdout


def test_job_plan_tables_match_the_per_read_definitions():
    """job.JobPlan: window counts / BN groups / valid samples against the plain per-read definitions
    (dataset.py:435-445 windows, :374-377 batches), packing order = descending window count."""
    from nanopore_rnn_b200 import engine, job
    rng = np.random.default_rng(3)
    lengths = np.concatenate([rng.integers(1, 5000, 40), [100, 101, 99, 300, 301, 1]])
    for step, seq, start in ((100, 100, 0), (300, 100, 0), (30, 100, 0), (100, 100, 7)):
        counts = job.window_counts(lengths, step, start)
        brute_valid = 0
        for n, c in zip(lengths, counts):
            _, lens = dataset.windows_from_signal(np.zeros(n), seq, step, start)
            assert len(lens) == c
            brute_valid += int(lens.sum())
        plan = job.JobPlan(lengths, seq, step, start, 10, 1 << 30, 1)
        assert plan.valid_samples == brute_valid and plan.n_windows == counts.sum()
        assert np.all(np.diff(plan.counts) <= 0) and sorted(plan.order) == list(range(len(lengths)))
        assert np.array_equal(plan.group_off, engine.group_offsets(plan.counts, 10))
        assert np.array_equal(np.diff(plan.read_off), lengths[plan.order])
    assert np.array_equal(job.group_offsets_vec([0, 5, 0, 23], 10), engine.group_offsets([0, 5, 0, 23], 10))


def test_bench_read_set_is_exact_and_sharding_independent():
    sys.path.insert(0, ROOT)
    import bench
    from nanopore_rnn_b200 import job
    for step in (100, 300):
        lengths = bench.make_read_set(20000, step)
        assert job.window_counts(lengths, step).sum() == 20000 and lengths.min() >= 1
    assert np.array_equal(bench.synth_read(5, 1000), bench.synth_read(5, 1000))
    assert not np.array_equal(bench.synth_read(5, 1000), bench.synth_read(6, 1000))


def test_world_size_2_gloo_gather_consensus(tmp_path):
    """job.gather_consensus over gloo with CPU tensors: the N>1 gather path of the reads workload (bench.py, e2e)."""
    script = tmp_path / "g.py"
    script.write_text("""
import sys
sys.path.insert(0, %r)
import numpy as np, torch, torch.distributed as dist
import nanopore_rnn_b200
from nanopore_rnn_b200 import shard, job
dist.init_process_group("gloo")
r, w = dist.get_rank(), dist.get_world_size()
counts = [5, 3, 9, 1, 1, 7, 2, 4]
mine = shard.assign_reads(counts, w)[r]
seqs = [("ACGT"[i %% 4] * (i + 1) + str(i)).encode() for i in mine]          # read i -> its own byte string
order = list(reversed(range(len(mine))))                                        # any packing order
cons = torch.from_numpy(np.frombuffer(b"".join(seqs[k] for k in order), dtype=np.uint8).copy())
off = torch.from_numpy(np.concatenate([[0], np.cumsum([len(seqs[k]) for k in order])]).astype(np.int64))
ids = np.asarray([mine[k] for k in order])
out = job.gather_consensus(cons, off, len(mine), ids, len(counts), r, w, "cpu")
if r == 0:
    assert out == [("ACGT"[i %% 4] * (i + 1) + str(i)).encode() for i in range(len(counts))], out
    print("OK", len(out))
else:
    assert out is None
dist.barrier()
dist.destroy_process_group()
""" % ROOT)
    out = subprocess.run([sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node=2",
                          "--master-addr", "127.0.0.1", "--master-port", "29619", str(script)],
                         capture_output=True, text=True, timeout=240)
    assert out.returncode == 0, out.stderr[-2000:]
    assert "OK 8" in out.stdout


def test_list_dir_uses_the_reference_predicate(tmp_path):
    """nanotensor/utils.py:71-82: regular files whose LAST dot component equals ext (not endswith)."""
    from nanopore_rnn_b200.utils import list_dir
    for name in ("a.signal", "b.fullsignal", "notasignal", "c.signal.bak", "d.txt"):
        (tmp_path / name).write_text("1 2 3")
    (tmp_path / "dir.signal").mkdir()
    got = [os.path.basename(p) for p in list_dir(str(tmp_path), ext="signal")]
    assert got == ["a.signal"]
    assert len(list_dir(str(tmp_path))) == 5                      # ext "": every regular file, no directories


def test_module_main_runs_the_job_once(tmp_path):
    """`python -m nanopore_rnn_b200.run_nanotensor -c cfg` must call main() exactly once (round-1 bug: twice) and
    return the Usage exit code 2 of THAT run."""
    cfg = tmp_path / "bad.json"
    cfg.write_text(json.dumps({"train": False, "inference": False, "test": False, "network": [],
                               "CreateDataset": {"dataset": "FullSignalSequence"}}))
    env = dict(os.environ, PYTHONPATH=ROOT + os.pathsep + os.environ.get("PYTHONPATH", ""))
    p = subprocess.run([sys.executable, "-m", "nanopore_rnn_b200.run_nanotensor", "-c", str(cfg)],
                       capture_output=True, text=True, cwd=ROOT, env=env, timeout=300)
    assert p.returncode == 2, (p.returncode, p.stderr[-500:])
    assert p.stderr.count("Need to specify exactly one of") == 1, p.stderr
