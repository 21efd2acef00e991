"""GPU (-m gpu): the device-resident inference job (job.ReadJob: pack -> forward -> consensus with every intermediate in
HBM) against the host route through the same C ABI, and the sharded form (whole reads per rank by LPT,
results read for read identical to the one-rank run)."""
import numpy as np
import pytest

from nanopore_rnn_b200 import assembly, dataset, engine, job, shard, synthetic as syn

pytestmark = pytest.mark.gpu


def _model(T=100):
    net = syn.NETWORKS["signal_CRNN"]
    w = syn.random_weights(net, 1, 5, seed=21, readout_gain=8.0)
    return engine.Model(net, w, "CtcLoss", T, 1, 5, device="cuda:0")


def _reads(seed, lengths):
    rng = np.random.default_rng(seed)
    return [np.rint(rng.standard_normal(n) * 60 + 500 + 30 * np.sin(np.arange(n) * 0.01)).astype(np.int16) for n in lengths]


def _host_route(m, reads, order, step, batch):
    """read_signal's z-score + windows on the host (dataset.py:435-445), forward_host, host-buffer assembly."""
    xs, ls, counts = [], [], []
    for r in order:
        sig = reads[r].astype(np.float64)
        sig = (sig - sig.mean()) / sig.std()
        x, lens = dataset.windows_from_signal(sig, m.T, step, 0)
        xs.append(x)
        ls.append(lens)
        counts.append(len(x))
    x, lens = np.concatenate(xs), np.concatenate(ls)
    labels, llen, _ = m.forward_host(x, lens, engine.group_offsets(counts, batch))
    return labels, np.concatenate([[0], np.cumsum(counts)]).astype(np.int64)


@pytest.mark.parametrize("step", [100, 300])
def test_device_chain_equals_host_route_on_ragged_reads(step):
    m = _model()
    lengths = [2311, 100, 37, 5000, 999, 1000, 1001, 64, 12345]
    reads = _reads(5, lengths)
    j = job.ReadJob(m, step=step, start_index=0, batch_size=20)
    plan, host = j.stage_raw(reads)
    assert np.all(np.diff(plan.counts) <= 0)                       # length-sorted packing
    cons, off = j.run_host(host, plan, keep_labels=True)
    chain_labels = j.last["labels"].cpu().numpy()
    chain_lens = j.last["lens"].cpu().numpy()
    seqs = j.to_original_order(cons, off, plan.order)
    # (1) the chain's consensus is the assembly of the chain's own labels (host-buffer entry of the same library)
    want = assembly.assemble_reads(chain_labels, plan.win_off)
    assert [s.decode() for s in seqs] == [want[int(np.nonzero(plan.order == r)[0][0])] for r in range(len(reads))]
    # (2) the chain's windows and labels equal the host route's (z-score in float64, same BN groups)
    host_labels, win_off = _host_route(m, reads, plan.order, step, 20)
    assert np.array_equal(win_off, plan.win_off)
    exp_lens = np.concatenate([dataset.windows_from_signal(np.zeros(lengths[r]), m.T, step, 0)[1] for r in plan.order])
    assert np.array_equal(chain_lens, exp_lens)
    bad = np.nonzero((chain_labels != host_labels).any(axis=1))[0]
    assert len(bad) <= max(1, len(chain_labels) // 1000), bad[:10]  # a float32 z-score rounding may flip a near tie
    if len(bad) == 0:
        assert want == assembly.assemble_reads(host_labels, win_off)
    assert sum(len(s) for s in seqs) > 50                           # not degenerate


def test_two_rank_split_reproduces_the_one_rank_reads():
    """shard.assign_reads over 2 ranks (run back to back on this one GPU): every read's labels and consensus are those
    of the one-rank job (BN groups never span reads, windows are independent)."""
    m = _model()
    lengths = syn.lognormal_read_lengths(24, seed=20, mu=np.log(3000.0), sigma=0.8, lo=150, hi=20000)
    reads = _reads(9, lengths)
    j = job.ReadJob(m, step=100, start_index=0, batch_size=20)

    def run(ids):
        plan, host = j.stage_raw([reads[i] for i in ids])
        cons, off = j.run_host(host, plan, keep_labels=True)
        labels = j.last["labels"].cpu().numpy()
        seqs = j.to_original_order(cons, off, plan.order)
        per_read = {}
        for pos, r in enumerate(plan.order):
            per_read[ids[int(r)]] = (labels[plan.win_off[pos]:plan.win_off[pos + 1]], seqs[int(r)])
        return per_read

    one = run(list(range(len(reads))))
    counts = job.window_counts(lengths, 100)
    parts = shard.assign_reads(counts, 2)
    assert abs(sum(counts[parts[0]]) - sum(counts[parts[1]])) <= counts.max()
    two = {}
    for ids in parts:
        two.update(run(ids))
    assert sorted(two) == sorted(one)
    flipped_reads = 0
    for r in one:
        la, lb = one[r][0], two[r][0]
        assert la.shape == lb.shape
        bad = int((la != lb).any(axis=1).sum())
        if bad:
            flipped_reads += 1
            assert bad <= max(1, len(la) // 500), (r, bad)          # near-tie flips from the BN block-sum rounding only
        else:
            assert one[r][1] == two[r][1]
    assert flipped_reads <= 1


def test_signal_files_through_the_chain(tmp_path):
    """.signal text -> ntb_parse_signal -> chain, against the host parse + host route."""
    m = _model()
    reads = _reads(3, [1500, 730, 2600])
    paths = []
    for i, r in enumerate(reads):
        p = tmp_path / ("r%d.signal" % i)
        p.write_text(" ".join(str(int(v)) for v in r) + "\n")
        paths.append(str(p))
    j = job.ReadJob(m, step=100, start_index=0, batch_size=20)
    seqs = j.run_signal_files(paths)
    plan, host = j.stage_raw(reads)
    cons, off = j.run_host(host, plan)
    assert seqs == [s.decode() for s in j.to_original_order(cons, off, plan.order)]
    assert all(len(s) > 5 for s in seqs)
