"""not gpu: the N > 1 host logic on CPU with world_size-2 gloo.  Each rank histograms its own
block-aligned shard, ONE all-reduce gives the global histogram, every rank normalises it with the
product's host adjust_freqs -> the same model everywhere (no broadcast), equal to the reference's
single-process model (golden vector)."""
import os
import socket
import sys

import numpy as np
import pytest

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(HERE)


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _worker(rank, world, port, out_dir):
    sys.path.insert(0, ROOT)
    sys.path.insert(0, HERE)
    import torch
    import torch.distributed as dist

    import ans_large_alphabet_b200 as P
    from ans_large_alphabet_b200 import sharded

    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    host = P.host_api()
    res = {}
    for name in ("fold1_zipf20", "msb_zipf20", "int_geom"):
        g = np.load(os.path.join(HERE, "golden", name + ".npz"))
        method, f, data = int(g["method"]), int(g["fidelity"]), g["data"]
        bs = 1024
        lo, hi = sharded.shard_range(data.size, bs, rank, world)
        mine = data[lo:hi]
        if method == P.METHOD_INT:
            mx = sharded.reduce_max(int(mine.max()) if mine.size else 0, "cpu", dist.group.WORLD)
            local = np.bincount(mine, minlength=mx + 1).astype(np.int64)
            # the one-pass variant the GPU path uses: optimistic counts + one-hot maxima in ONE all-reduce
            K = sharded.OPT_BINS
            buf = np.zeros(K + world, dtype=np.int64)
            inside = mine[mine < K]
            buf[:K] = np.bincount(inside, minlength=K)[:K]
            buf[K + rank] = int(mine.max()) if mine.size else 0
            counts, gmax = sharded.reduce_bins_and_max(torch.from_numpy(buf), K, dist.group.WORLD)
            assert gmax == mx, (gmax, mx)
            res[name + "_onepass"] = counts[: gmax + 1].copy() if gmax < K else np.zeros(0, np.uint64)
        else:
            nb = sharded.alphabet_bins(method, f)
            local = np.zeros(nb, dtype=np.int64)
            uniq, cnt = np.unique(mine, return_counts=True)
            for v, c in zip(uniq, cnt):
                s = host.msb_map(v)[0] if method == P.METHOD_MSB else host.fold_map(f, v)[0]
                local[s] += c
        bins = sharded.reduce_histogram(torch.from_numpy(local), dist.group.WORLD)
        largest = int(np.nonzero(bins)[0][-1])
        nfreqs = host.adjust_freqs(bins, largest, method != P.METHOD_INT)
        res[name + "_nfreqs"] = nfreqs
        res[name + "_range"] = np.array([lo, hi])
        res[name + "_total"] = np.array([int(bins.sum())])
    np.savez(os.path.join(out_dir, f"rank{rank}.npz"), **res)
    dist.barrier()
    dist.destroy_process_group()


def test_shard_ranges_partition_the_blocks():
    from ans_large_alphabet_b200.sharded import shard_range

    for n, bs, world in [(10_000_000, 8192, 8), (8192 * 5 + 3, 8192, 4), (100, 8192, 2), (1 << 30, 8192, 8), (70_001, 128, 3)]:
        prev = 0
        for r in range(world):
            lo, hi = shard_range(n, bs, r, world)
            assert lo == prev and lo % bs == 0 and (hi % bs == 0 or hi == n)
            prev = hi
        assert prev == n


def test_two_rank_global_model_equals_reference(tmp_path):
    import torch.multiprocessing as mp

    port = _free_port()
    mp.spawn(_worker, args=(2, port, str(tmp_path)), nprocs=2, join=True)
    r0, r1 = np.load(tmp_path / "rank0.npz"), np.load(tmp_path / "rank1.npz")
    for name in ("fold1_zipf20", "msb_zipf20", "int_geom"):
        g = np.load(os.path.join(HERE, "golden", name + ".npz"))
        assert np.array_equal(r0[name + "_nfreqs"], r1[name + "_nfreqs"]), "ranks disagree on the model"
        assert np.array_equal(r0[name + "_nfreqs"], g["nfreqs"]), "global model differs from the reference's"
        assert int(r0[name + "_total"][0]) == g["data"].size
        assert r0[name + "_range"][1] == r1[name + "_range"][0] and r1[name + "_range"][1] == g["data"].size
    # plain ANS, one pass + one all-reduce: the same global counts on both ranks, equal to a histogram of the whole input
    g = np.load(os.path.join(HERE, "golden", "int_geom.npz"))
    want = np.bincount(g["data"]).astype(np.uint64)
    assert np.array_equal(r0["int_geom_onepass"], want) and np.array_equal(r1["int_geom_onepass"], want)


# ---- -m gpu: the library's own NCCL exchange, two ranks (needs two GPUs; the 1-GPU box skips it) -------------------
def _gpu_worker(rank, world, port, out_dir, exchange):
    os.environ["B200ANS_NO_IPC"] = "0" if exchange == "peer_memory" else "1"
    sys.path.insert(0, ROOT)
    sys.path.insert(0, HERE)
    import torch
    import torch.distributed as dist

    import ans_large_alphabet_b200 as P
    from ans_large_alphabet_b200 import sharded

    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    torch.cuda.set_device(rank)
    dist.init_process_group("gloo", rank=rank, world_size=world)  # only carries the NCCL unique id
    codec = P.Codec(device=rank, block_syms=1024)
    sharded.init_comm(codec, dist.group.WORLD)
    res = {}
    for name in ("fold1_zipf20", "msb_zipf20", "int_geom", "rfold2_zipf20"):
        g = np.load(os.path.join(HERE, "golden", name + ".npz"))
        method, f, data = int(g["method"]), int(g["fidelity"]), g["data"]
        lo, hi = sharded.shard_range(data.size, 1024, rank, world)
        if name == "msb_zipf20":  # rank 1's shard is empty: it takes part in the reduction, writes an empty container
            lo, hi = (0, data.size) if rank == 0 else (data.size, data.size)
        mine = np.ascontiguousarray(data[lo:hi])
        d_in = torch.from_numpy(mine.view(np.int32)).cuda() if mine.size else torch.zeros(4, dtype=torch.int32, device="cuda")
        d_blob = torch.empty((codec.bound(method, max(mine.size, 1)) + 15) // 16 * 16, dtype=torch.uint8, device="cuda")
        torch.cuda.synchronize()
        nb = codec.compress_sharded_device(method, f, d_blob.data_ptr(), d_blob.numel(), d_in.data_ptr(), mine.size)
        blob = d_blob[:nb].cpu().numpy()
        frame, nfreqs = codec.model_nfreqs()
        res[name + "_nfreqs"] = nfreqs
        res[name + "_prelude"] = codec.model_prelude()
        res[name + "_blob"] = blob
        res[name + "_range"] = np.array([lo, hi])
        back = codec.decompress(method, f, blob, mine.size)
        res[name + "_ok"] = np.array([int(np.array_equal(back, mine))])
    # fold-1, msb and plain ANS exchange 8 KiB .. 24 KiB: over peer memory when it is set up; ANSrfold's raw bins: NCCL
    res["peer_exchanges"] = np.array([codec.stats().peer_exchanges])
    np.savez(os.path.join(out_dir, f"gpu_rank{rank}.npz"), **res)
    dist.barrier()  # nobody unmaps an exchange buffer a peer may still be writing into
    codec.close()
    dist.barrier()
    dist.destroy_process_group()


@pytest.mark.gpu
@pytest.mark.parametrize("exchange", ["peer_memory", "nccl"])  # b200ans_comm_ipc_* (one kernel over NVLink) / ncclAllReduce
def test_two_gpu_sharded_compress_uses_the_global_model(tmp_path, exchange):
    import torch

    if torch.cuda.device_count() < 2:
        pytest.skip("needs two GPUs")
    import torch.multiprocessing as mp

    import _container
    from _libs import have_reference, oracle, reference

    cpu = reference() if have_reference() else oracle()
    port = _free_port()
    mp.spawn(_gpu_worker, args=(2, port, str(tmp_path), exchange), nprocs=2, join=True)
    r = [np.load(tmp_path / "gpu_rank0.npz"), np.load(tmp_path / "gpu_rank1.npz")]
    for k in (0, 1):
        n_peer = int(r[k]["peer_exchanges"][0])
        assert (n_peer >= 3) if exchange == "peer_memory" else (n_peer == 0), n_peer
    for name in ("fold1_zipf20", "msb_zipf20", "int_geom", "rfold2_zipf20"):
        g = np.load(os.path.join(HERE, "golden", name + ".npz"))
        method, f, data = int(g["method"]), int(g["fidelity"]), g["data"]
        assert np.array_equal(r[0][name + "_nfreqs"], g["nfreqs"]) and np.array_equal(r[1][name + "_nfreqs"], g["nfreqs"]), name
        assert _container.prelude_equal(r[0][name + "_prelude"], g["prelude"]) and np.array_equal(r[0][name + "_prelude"], r[1][name + "_prelude"])
        assert int(r[0][name + "_ok"][0]) == 1 and int(r[1][name + "_ok"][0]) == 1
        model = cpu.model_create(method, f, data)
        for k in (0, 1):
            lo, hi = (int(x) for x in r[k][name + "_range"])
            if hi == lo:
                assert r[k][name + "_blob"].size == 64
                continue
            c = _container.parse(r[k][name + "_blob"])
            for b, payload in enumerate(c["blocks"]):
                want = model.encode_block(data[lo + b * 1024 : min(hi, lo + (b + 1) * 1024)])
                assert np.array_equal(payload, want), (name, k, b)
