"""Multi-GPU host logic: shard planning, the one exchange step and the ordered merge.

CPU: two gloo ranks, each holding the per-shard results the GPU would produce (taken from the oracle run
on the shard alone), must reassemble exactly the oracle's whole-file result.
GPU (-m gpu): the same decomposition on one device, shard after shard, through xsg_scan + xsg_result_rebase."""
import os
import socket

import numpy as np
import pytest

import gen_cases
import oracle_lib


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _corpus():
    import xsearch_b200 as xs
    p = xs.SynthParams.make(31, b"Sherlock", 30.0 / 4096)
    buf = xs.synth_host(4096 * 700, p)
    return buf, xs.chunk_bounds_host(buf, 200000)


def test_shard_chunks_partition():
    import xsearch_b200 as xs
    buf, bounds = _corpus()
    for world in (1, 2, 3, 4, 8, 64):
        sh = xs.shard_chunks(bounds, world)
        assert len(sh) == world and sh[0][0] == 0 and sh[-1][1] == bounds.size
        assert all(a[1] == b[0] for a, b in zip(sh, sh[1:])) and all(c0 <= c1 for c0, c1 in sh)
        if world <= 8:
            sizes = [int(bounds[c1 - 1]) - (int(bounds[c0 - 1]) if c0 else 0) for c0, c1 in sh if c1 > c0]
            assert max(sizes) - min(sizes) <= 2 * 260000  # balanced to about a chunk
    assert xs.shard_chunks(np.zeros(0, np.uint64), 2) == [(0, 0), (0, 0)]


def _rank_main(rank, world, port, out_dir):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank), WORLD_SIZE=str(world))
    import torch.distributed as dist
    import xsearch_b200 as xs
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        oracle, _ = oracle_lib.load(build=False)
        buf, bounds = _corpus()
        c0, c1 = xs.shard_chunks(bounds, world)[rank]
        b0 = int(bounds[c0 - 1]) if c0 else 0
        b1 = int(bounds[c1 - 1]) if c1 > c0 else b0
        shard, sb = buf[b0:b1], bounds[c0:c1] - np.uint64(b0)
        pat = b"Sherlock"
        # what this rank's GPU would hold after xsg_scan(offset_base = b0): shard-local results + byte base
        match = oracle.match_byte_offsets(shard, pat, sb) + np.uint64(b0)
        lidx = oracle.line_indices(shard, pat, sb)
        newlines = int((shard == 10).sum())
        line_base, bases, totals = xs.shard_exchange(newlines, [match.size, lidx.size])
        lidx = lidx + np.uint64(line_base)  # == xsg_result_rebase(res, 0, line_base)
        np.savez(os.path.join(out_dir, "rank%d.npz" % rank), match=match, lidx=lidx, bases=np.array(bases),
                 totals=np.array(totals))
    finally:
        dist.destroy_process_group()


@pytest.mark.parametrize("world", [2, 3])
def test_two_rank_exchange_and_merge_gloo(tmp_path, world):
    import torch.multiprocessing as mp
    port = _free_port()
    mp.spawn(_rank_main, args=(world, port, str(tmp_path)), nprocs=world, join=True)
    oracle, _ = oracle_lib.load(build=False)
    buf, bounds = _corpus()
    exp_match = oracle.match_byte_offsets(buf, b"Sherlock", bounds)
    exp_lidx = oracle.line_indices(buf, b"Sherlock", bounds)
    parts = [np.load(os.path.join(str(tmp_path), "rank%d.npz" % r)) for r in range(world)]
    assert np.array_equal(np.concatenate([p["match"] for p in parts]), exp_match)   # rank order == file order
    assert np.array_equal(np.concatenate([p["lidx"] for p in parts]), exp_lidx)
    off = 0
    for p in parts:  # the output base of a rank is the number of results before it
        assert int(p["bases"][0]) == off and int(p["totals"][0]) == exp_match.size and int(p["totals"][1]) == exp_lidx.size
        off += p["match"].size


@pytest.mark.gpu
def test_shards_on_one_device_equal_whole():
    import xsearch_b200 as xs
    oracle, _ = oracle_lib.load(build=False)
    buf, bounds = _corpus()
    kinds = xs.MATCH_BYTE_OFFSETS | xs.LINE_BYTE_OFFSETS | xs.LINE_INDICES | xs.LINES | xs.COUNT | xs.COUNT_LINES
    ctx = xs.Context(0)
    for pat in (b"Sherlock", b"the"):
        got = {k: [] for k in ("match", "loff", "lidx", "lst", "lln")}
        line_base = cnt = cl = 0
        for c0, c1 in xs.shard_chunks(bounds, 3):
            b0 = int(bounds[c0 - 1]) if c0 else 0
            corpus = ctx.corpus_from_host(buf[b0:int(bounds[c1 - 1])], bounds=bounds[c0:c1] - np.uint64(b0))
            res = ctx.scan(corpus, pat, kinds, offset_base=b0)
            res.rebase(0, line_base)             # the base the exchange would deliver
            line_base += res.newlines
            cnt += res.count(xs.COUNT)
            cl += res.count(xs.COUNT_LINES)
            got["match"].append(res.vector(xs.MATCH_BYTE_OFFSETS))
            got["loff"].append(res.vector(xs.LINE_BYTE_OFFSETS))
            got["lidx"].append(res.vector(xs.LINE_INDICES))
            st, ln = res.line_spans()
            got["lst"].append(st)
            got["lln"].append(ln)
            res.free()
            corpus.free()
        assert cnt == oracle.count(buf, pat, bounds, False) and cl == oracle.count(buf, pat, bounds, True)
        assert np.array_equal(np.concatenate(got["match"]), oracle.match_byte_offsets(buf, pat, bounds))
        assert np.array_equal(np.concatenate(got["loff"]), oracle.line_byte_offsets(buf, pat, bounds))
        assert np.array_equal(np.concatenate(got["lidx"]), oracle.line_indices(buf, pat, bounds))
        est, eln = oracle.lines(buf, pat, bounds)
        assert np.array_equal(np.concatenate(got["lst"]), est) and np.array_equal(np.concatenate(got["lln"]), eln)
    ctx.close()


@pytest.mark.gpu
def test_device_side_summary_and_table_rebase():
    """The host-free exchange: the scan's summary lives on the device, the fix-up kernel reads the gathered table."""
    import torch
    import xsearch_b200 as xs
    oracle, _ = oracle_lib.load(build=False)
    buf, bounds = _corpus()
    ctx = xs.Context(0)
    corpus = ctx.corpus_from_host(buf, bounds=bounds)
    kinds = xs.COUNT | xs.COUNT_LINES | xs.LINE_INDICES | xs.LINES | xs.MATCH_BYTE_OFFSETS
    res = ctx.scan(corpus, b"Sherlock", kinds)
    torch.cuda.synchronize()
    summ = ctx.summary_tensor().cpu().numpy()
    assert list(summ) == [res.newlines, res.count(xs.COUNT), res.count(xs.COUNT_LINES), res.count(xs.LINES)]
    assert summ[0] == int((buf == 10).sum())
    # pretend to be rank 2 of 3: ranks 0 and 1 hold 1000 and 234 newlines
    table = torch.tensor([[1000, 1, 1, 1], [234, 2, 2, 2], [int(summ[0]), 3, 3, 3]], dtype=torch.int64, device="cuda")
    res.rebase_from_table(table, 2)
    torch.cuda.synchronize()
    assert np.array_equal(res.vector(xs.LINE_INDICES), oracle.line_indices(buf, b"Sherlock", bounds) + np.uint64(1234))
    assert np.array_equal(res.vector(xs.MATCH_BYTE_OFFSETS), oracle.match_byte_offsets(buf, b"Sherlock", bounds))
    res.free()
    # count-only and direct paths fill the summary as well
    res = ctx.scan(corpus, b"Sherlock", xs.COUNT)
    torch.cuda.synchronize()
    assert int(ctx.summary_tensor()[1].item()) == res.count(xs.COUNT) == oracle.count(buf, b"Sherlock", bounds, False)
    res.free()
    corpus.free()
    ctx.close()


def _peer_rank_main(rank, world, port, out_dir):
    """One rank of the NVLink peer-memory exchange: scan the shard, push the summary into every rank's table,
    fix the line indices up on the device (xsg_exchange_*), three rounds with three patterns."""
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank), WORLD_SIZE=str(world),
                      LOCAL_RANK=str(rank))
    import torch
    import torch.distributed as dist
    import xsearch_b200 as xs
    torch.cuda.set_device(rank)
    dist.init_process_group("nccl", rank=rank, world_size=world, device_id=torch.device("cuda", rank))
    try:
        buf, bounds = _corpus()
        c0, c1 = xs.shard_chunks(bounds, world)[rank]
        b0 = int(bounds[c0 - 1]) if c0 else 0
        ctx = xs.Context(rank)
        corpus = ctx.corpus_from_host(buf[b0:int(bounds[c1 - 1])], bounds=bounds[c0:c1] - np.uint64(b0))
        ex = xs.Exchange(ctx)
        out = {}
        for i, pat in enumerate((b"Sherlock", b"the", b"zzzz-not-there")):
            if rank == i % world:
                torch.cuda._sleep(200_000_000)   # one rank is late: the others must wait on the device
            # rounds alternate between the two ways to publish: from inside the scan, and with a launch of its own
            res = ctx.scan(corpus, pat, xs.MATCH_BYTE_OFFSETS | xs.LINE_INDICES | xs.COUNT_LINES, offset_base=b0,
                           publish=(i % 2 == 0))
            if i % 2:
                ex.push(res)
            ex.rebase(res)
            out["m%d" % i] = res.vector(xs.MATCH_BYTE_OFFSETS)
            out["l%d" % i] = res.vector(xs.LINE_INDICES)
            out["t%d" % i] = ex.totals()
            res.free()
        np.savez(os.path.join(out_dir, "peer%d.npz" % rank), **out)
        ex.close()
        corpus.free()
        ctx.close()
    finally:
        dist.destroy_process_group()


def _check_peer_outputs(out_dir, world):
    oracle, _ = oracle_lib.load(build=False)
    buf, bounds = _corpus()
    parts = [np.load(os.path.join(out_dir, "peer%d.npz" % r)) for r in range(world)]
    for i, pat in enumerate((b"Sherlock", b"the", b"zzzz-not-there")):
        exp_m = oracle.match_byte_offsets(buf, pat, bounds)
        exp_l = oracle.line_indices(buf, pat, bounds)
        assert np.array_equal(np.concatenate([p["m%d" % i] for p in parts]), exp_m), pat
        assert np.array_equal(np.concatenate([p["l%d" % i] for p in parts]), exp_l), pat
        for p in parts:  # every rank ends up with the same gathered table
            t = p["t%d" % i]
            assert t.shape == (world, 4) and int(t[:, 0].sum()) == int((buf == 10).sum())
            assert int(t[:, 1].sum()) == exp_m.size and int(t[:, 2].sum()) == exp_l.size


@pytest.mark.gpu
def test_peer_exchange_single_rank(tmp_path):
    """world = 1: the exchange degenerates to a rank writing into its own table (kernels, rounds, parities)."""
    import torch.multiprocessing as mp
    mp.spawn(_peer_rank_main, args=(1, _free_port(), str(tmp_path)), nprocs=1, join=True)
    _check_peer_outputs(str(tmp_path), 1)


@pytest.mark.gpu
def test_peer_exchange_two_gpus(tmp_path):
    """Two processes, two GPUs: summaries travel as peer stores over NVLink, no collective call per round."""
    import torch
    if torch.cuda.device_count() < 2:
        pytest.skip("needs two GPUs (run with gpurun --gpus 2)")
    import torch.multiprocessing as mp
    mp.spawn(_peer_rank_main, args=(2, _free_port(), str(tmp_path)), nprocs=2, join=True)
    _check_peer_outputs(str(tmp_path), 2)


# ---- one process, several devices (xsg_init_multi) ------------------------------------------------------------------
def _n_gpus():
    import torch
    return torch.cuda.device_count()


@pytest.mark.gpu
def test_search_host_multi_single_device_list():
    """xsg_search_host_multi with a list of ONE device is xsg_search_host (same code path, no local line indices)."""
    import xsearch_b200 as xs
    oracle, _ = oracle_lib.load(build=False)
    buf, bounds = _corpus()
    mc = xs.MultiContext([0])
    kinds = xs.COUNT_LINES | xs.LINE_INDICES | xs.MATCH_BYTE_OFFSETS | xs.LINES
    os.environ["XSG_SEGMENT_MB"] = "1"
    try:
        res = mc.search_host(buf, None, b"Sherlock", kinds, bounds=bounds, line_index_base=7)
    finally:
        del os.environ["XSG_SEGMENT_MB"]
    assert np.array_equal(res.vector(xs.LINE_INDICES), oracle.line_indices(buf, b"Sherlock", bounds) + np.uint64(7))
    assert np.array_equal(res.vector(xs.MATCH_BYTE_OFFSETS), oracle.match_byte_offsets(buf, b"Sherlock", bounds))
    res.free()
    mc.close()


@pytest.mark.gpu
def test_search_host_multi_devices():
    """ONE host buffer, segments dealt round-robin to all devices of the box, ONE result in file order with global
    offsets and line indices (bases added on the device, no collective): bit-identical to the oracle on the whole
    buffer, for several segment sizes (more, as many and fewer segments than devices)."""
    n = _n_gpus()
    if n < 2:
        pytest.skip("needs two GPUs (run with gpurun --gpus 2)")
    import xsearch_b200 as xs
    oracle, _ = oracle_lib.load(build=False)
    buf, bounds = _corpus()
    mc = xs.MultiContext(list(range(min(n, 4))))
    kinds = xs.COUNT | xs.COUNT_LINES | xs.LINE_INDICES | xs.LINE_BYTE_OFFSETS | xs.MATCH_BYTE_OFFSETS | xs.LINES
    for seg_mb, pat in (("1", b"Sherlock"), ("1", b"the"), ("2", b"that"), ("64", b"Sherlock")):
        os.environ["XSG_SEGMENT_MB"] = seg_mb
        try:
            res = mc.search_host(buf, None, pat, kinds, bounds=bounds, offset_base=1000, line_index_base=5)
        finally:
            del os.environ["XSG_SEGMENT_MB"]
        assert res.count(xs.COUNT) == oracle.count(buf, pat, bounds, False)
        assert res.count(xs.COUNT_LINES) == oracle.count(buf, pat, bounds, True)
        assert np.array_equal(res.vector(xs.MATCH_BYTE_OFFSETS), oracle.match_byte_offsets(buf, pat, bounds) + np.uint64(1000))
        assert np.array_equal(res.vector(xs.LINE_BYTE_OFFSETS), oracle.line_byte_offsets(buf, pat, bounds) + np.uint64(1000))
        assert np.array_equal(res.vector(xs.LINE_INDICES), oracle.line_indices(buf, pat, bounds) + np.uint64(5))
        st, ln = res.line_spans()
        es, el = oracle.lines(buf, pat, bounds)
        assert np.array_equal(st, es + np.uint64(1000)) and np.array_equal(ln, el)
        assert res.newlines == int((buf == 10).sum())
        res.free()
    mc.close()


@pytest.mark.gpu
def test_local_exchange_devices():
    """Device-resident shards on several devices of ONE process: the exchanges are connected through peer access
    (xsg_exchange_connect_local, no IPC handles); publishing from inside the scan and with xsg_exchange_push, a
    streamed multi-part result included; publishing twice without waiting is refused."""
    n = _n_gpus()
    if n < 2:
        pytest.skip("needs two GPUs (run with gpurun --gpus 2)")
    import xsearch_b200 as xs
    oracle, _ = oracle_lib.load(build=False)
    buf, bounds = _corpus()
    world = min(n, 4)
    mc = xs.MultiContext(list(range(world)))
    exs = mc.exchanges()
    shards = xs.shard_chunks(bounds, world)
    corp, base, sb = [], [], []
    for r, (c0, c1) in enumerate(shards):
        b0 = int(bounds[c0 - 1]) if c0 else 0
        b1 = int(bounds[c1 - 1]) if c1 > c0 else b0
        base.append(b0)
        sb.append(bounds[c0:c1] - np.uint64(b0))
        corp.append(mc[r].corpus_from_host(buf[b0:b1], bounds=sb[-1]))
    kinds = xs.MATCH_BYTE_OFFSETS | xs.LINE_INDICES | xs.COUNT_LINES
    for rnd, pat in enumerate((b"Sherlock", b"the", b"zzzz-not-there")):
        exp_m, exp_l = oracle.match_byte_offsets(buf, pat, bounds), oracle.line_indices(buf, pat, bounds)
        res = []
        for r in range(world):  # every rank publishes ...
            if rnd == 1:  # ... a streamed (multi-part) result with a launch of its own
                c0, c1 = shards[r]
                b1 = int(bounds[c1 - 1]) if c1 > c0 else base[r]
                os.environ["XSG_SEGMENT_MB"] = "1"
                try:
                    q = mc[r].search_host(buf[base[r]:b1], None, pat, kinds, bounds=sb[r], offset_base=base[r])
                finally:
                    del os.environ["XSG_SEGMENT_MB"]
                exs[r].push(q)
            else:  # ... from inside the scan
                q = mc[r].scan(corp[r], pat, kinds, offset_base=base[r], publish=True)
            res.append(q)
        if rnd == 0:  # a second round before the first was waited for would overrun the slot parities
            with pytest.raises(xs.XsgError):
                exs[0].push(res[0])
        for r in range(world):  # ... then every rank waits for all and fixes its line indices up
            exs[r].rebase(res[r])
        assert np.array_equal(np.concatenate([q.vector(xs.MATCH_BYTE_OFFSETS) for q in res]), exp_m), pat
        assert np.array_equal(np.concatenate([q.vector(xs.LINE_INDICES) for q in res]), exp_l), pat
        t = exs[world - 1].totals()
        assert int(t[:, 0].sum()) == int((buf == 10).sum()) and int(t[:, 1].sum()) == exp_m.size and int(t[:, 2].sum()) == exp_l.size
        for r in range(world):
            exs[r].check()
            res[r].free()
    for e in exs:
        e.close()
    for c in corp:
        c.free()
    mc.close()
