#!/usr/bin/env python
"""bench.py -- headline benchmark of the b200ans hot path (see DESIGN.md "Measurement").

One "step" = one semi-static ENCODE pass (histogram K1 -> host normalise -> block encoder K2 ->
size scan + compaction) followed by one DECODE pass (prelude -> tables -> block decoder K3) over
one batch of synthetic uint32 symbols.  metric = GB/s of uint32 input, i.e. (4n + 4n) bytes per
step / step time (the mean over the encode and the decode pass); the two passes are also
reported separately.

  value  : inputs resident in HBM when the timed region starts (device entry points)
  e2e    : the same step through the host-buffer entry points b200ans_compress /
           b200ans_decompress (what methods.hpp-shaped adaptors call) with pinned HOST buffers,
           H2D / D2H copies inside the timed region
  --impl reference : the reference's own single-threaded CPU coder (oracle/_ref/libansref.so,
           the unmodified headers; falls back to the C port) on a bounded sample per step

The default workload is BASELINE.json's config 5 as ONE FIXED JOB: semi-static ANSfold-1 over 2^32 symbols (the
configuration the north star's >= 7x-on-8-GPUs target is stated on; it fits one B200), split into N contiguous shards
of 2^32 / N symbols -> "scaling": "strong".  The same line carries `north_star_2p30`: the 2^30-symbol single-GPU
targets (ANSfold-1, ANSfold-2, ANS) kernel by kernel, and `parity`: blocks of the step's container compared byte for
byte with the reference coder (oracle/_ref) run on those blocks with the same global model, outside the timed region.

python bench.py [--gpus N] [--steps K] [--warmup W] [--workload c5|c2|c1|c3|c3r|fold1|...] [--impl ours|reference]
"""
import argparse
import json
import math
import os
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))

M_INT, M_FOLD, M_MSB, M_RFOLD = 0, 1, 2, 3

# BASELINE.json configs (SURVEY.md section 8 shorthand) + the north_star's 2^30 target sizes.
WORKLOADS = {
    # name: (method, fidelity, generator, n, description)
    "c1": (M_FOLD, 1, ("zipf", 1 << 20, 1.0, 0), 10_000_000, "C1 ANSfold-1, 1e7-symbol Zipf(1.0, 2^20)"),
    "c2": (M_INT, 0, ("geom", 0.01), 100_000_000, "C2 ANS (ans_int), 1e8-symbol geometric(p=0.01)"),
    "c3": (M_FOLD, 2, ("zipf", 1 << 24, 1.0, 0), 1_000_000_000, "C3 ANSfold-2, 1e9-symbol Zipf(1.0, 2^24)"),
    "c3r": (M_RFOLD, 2, ("zipf", 1 << 24, 1.0, 0), 1_000_000_000, "C3 ANSrfold-2, 1e9-symbol Zipf(1.0, 2^24)"),
    "c5": (M_FOLD, 1, ("zipf", 1 << 20, 1.2, 1), 1 << 32, "C5 semi-static ANSfold-1, ONE 2^32-symbol BWT-MTF-style job (Zipf(1.2, 2^20)-1), split over the GPUs"),
    "c5w": (M_FOLD, 1, ("zipf", 1 << 20, 1.2, 1), 1 << 30, "C5-style ANSfold-1, 2^30-symbol Zipf(1.2, 2^20)-1 per GPU (weak scaling)"),
    "fold1": (M_FOLD, 1, ("zipf", 1 << 20, 1.0, 0), 1 << 30, "ANSfold-1, 2^30-symbol Zipf(1.0, 2^20)"),
    "fold2": (M_FOLD, 2, ("zipf", 1 << 20, 1.0, 0), 1 << 30, "ANSfold-2, 2^30-symbol Zipf(1.0, 2^20)"),
    "msb": (M_MSB, 0, ("zipf", 1 << 20, 1.0, 0), 1 << 30, "ANSmsb, 2^30-symbol Zipf(1.0, 2^20)"),
    "ans30": (M_INT, 0, ("geom", 0.01), 1 << 30, "ANS (ans_int), 2^30-symbol geometric(p=0.01)"),
    "c4": (M_FOLD, 1, ("zipf", 1 << 20, 1.0, 0), 1 << 27, "C4 pseudo-adaptive ANSfold-1 (one model per 8192 symbols), 2^27-symbol Zipf(1.0, 2^20) per GPU"),
}
# pseudo-adaptive workloads: symbols per model (every block has its own histogram / model / prelude; no collective)
PA_MODEL_SYMS = {"c4": 8192}
# Fixed-size jobs: the symbol count is the JOB's, every rank takes 1/N of it ("scaling": "strong").  Everything else
# keeps the symbol count per GPU ("weak").
STRONG = {"c5"}
DEFAULT_WORKLOAD = "c5"  # BASELINE.json configs[4]: the largest single-GPU configuration, the one the targets are stated on
# Symbols per block (a container parameter).  8192 is the library default; the 1e8-symbol C2 input has only 12 207 such
# blocks x 4 rANS chains = 48 828 independent chains for 148 SMs, so it runs with 4096 (twice the chains; the size
# overhead against the reference's whole stream, reported in the line, stays below the 1 % budget).
BLOCK_SYMS = {"c2": 4096, "c1": 4096}
METRIC = "encode+decode GB/s of uint32 input"
UNIT = "GB/s"


def method_label(method, f):
    return {M_INT: "ANS", M_MSB: "ANSmsb", M_FOLD: f"ANSfold-{f}", M_RFOLD: f"ANSrfold-{f}"}[method]


# ------------------------------------------------------------------------------------------------
# synthetic inputs
# ------------------------------------------------------------------------------------------------
def generate_device(gen, n, device, seed):
    """Synthetic symbols generated on the GPU (torch is plumbing here): same distributions as
    generate_inputs.cpp:97-117, different random stream."""
    import torch

    g = torch.Generator(device=device)
    g.manual_seed(1234 + seed)
    out = torch.empty(n, dtype=torch.int32, device=device)
    chunk = 1 << 26
    if gen[0] == "geom":
        p = gen[1]
        inv = 1.0 / math.log1p(-p)
        for lo in range(0, n, chunk):
            m = min(chunk, n - lo)
            u = torch.rand(m, device=device, generator=g, dtype=torch.float64)
            out[lo : lo + m] = torch.floor(torch.log1p(-u) * inv).to(torch.int32)
    elif gen[0] == "zipf":
        _, alphabet, q, sub = gen
        w = 1.0 / torch.arange(1, alphabet + 1, device=device, dtype=torch.float64) ** q
        cdf = torch.cumsum(w, 0)
        cdf /= cdf[-1].clone()
        for lo in range(0, n, chunk):
            m = min(chunk, n - lo)
            u = torch.rand(m, device=device, generator=g, dtype=torch.float64)
            r = torch.searchsorted(cdf, u).clamp_(max=alphabet - 1)
            out[lo : lo + m] = (r + (1 - sub)).to(torch.int32)
    else:
        raise KeyError(gen)
    return out


def generate_host(gen, n, seed):
    from _inputs import geometric, zipf

    if gen[0] == "geom":
        return geometric(n, gen[1], seed)
    _, alphabet, q, sub = gen
    return zipf(n, alphabet, q, seed, offset=1 - sub)


# ------------------------------------------------------------------------------------------------
# clocks sampler (B200_PROFILING.md "clocks DURING the timed region")
# ------------------------------------------------------------------------------------------------
class ClockSampler:
    Q = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
         "clocks_event_reasons.sw_power_cap")

    def __init__(self, index):
        self.index, self.rows, self.proc = index, [], None

    def start(self):
        try:
            self.proc = subprocess.Popen(
                ["nvidia-smi", f"--id={self.index}", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits", "-lms", "100"],
                stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            threading.Thread(target=self._pump, daemon=True).start()
        except Exception:
            self.proc = None

    def _pump(self):
        for line in self.proc.stdout:
            self.rows.append((time.time(), line.strip()))

    def stop(self, t0, t1):
        if not self.proc:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        time.sleep(0.15)
        self.proc.terminate()
        sm, smax, reasons = [], None, set()
        for ts, line in self.rows:
            parts = [p.strip() for p in line.split(",")]
            if len(parts) < 9:
                continue
            try:
                clk, mx = float(parts[1]), float(parts[2])
            except ValueError:
                continue
            smax = mx
            if t0 - 0.05 <= ts <= t1 + 0.15:
                sm.append(clk)
                for name, val in zip(("hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"), parts[5:9]):
                    if val.lower().startswith("active"):
                        reasons.add(name)
        if not sm:  # timed region shorter than the sampling period: take everything we saw
            for ts, line in self.rows:
                parts = [p.strip() for p in line.split(",")]
                try:
                    sm.append(float(parts[1]))
                except Exception:
                    pass
        return {"sm_mhz": float(np.median(sm)) if sm else None, "sm_max_mhz": smax, "reasons": sorted(reasons),
                "samples": len(sm)}


def measured_peak():
    try:
        with open(os.path.join(ROOT, "MEASURED_PEAKS.json")) as fh:
            return float(json.load(fh)["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs)"
    except Exception:
        return 6650.0, "fallback (B200_PROFILING.md 6.65 TB/s)"


# ------------------------------------------------------------------------------------------------
# CPU baseline: the reference's own coder on this box's host cores (test infrastructure, the one
# place besides tests/ and smoke() that executes anything under oracle/)
# ------------------------------------------------------------------------------------------------
def cpu_codec():
    from _libs import have_reference, oracle, reference

    if have_reference():
        return reference(), "reference"
    return oracle(), "port"


def cpu_time_sample(method, f, sample, runs, block=0):
    codec, kind = cpu_codec()
    enc_ns, dec_ns, nbytes = codec.time(method, f, sample, runs, block)
    gbs = 8.0 * sample.size / (enc_ns + dec_ns)  # (4n + 4n) bytes / (t_enc + t_dec); bytes/ns == GB/s
    return {
        "value": gbs, "unit": UNIT, "cores": 1, "kind": kind,
        "sample": f"first {sample.size} symbols of the workload, "
                  + (f"{method_label(method, f)} block by block ({block} symbols each)" if block else f"whole-stream {method_label(method, f)}")
                  + f", min of {runs} runs",
        "encode_gbs": 4.0 * sample.size / enc_ns, "decode_gbs": 4.0 * sample.size / dec_ns,
        "encode_ns_per_symbol": enc_ns / sample.size, "decode_ns_per_symbol": dec_ns / sample.size,
        "bits_per_symbol": 8.0 * nbytes / sample.size, "host_cores_total": os.cpu_count(),
        "note": "the reference is single-threaded (no threading anywhere in its sources): 1 of %d cores" % (os.cpu_count() or 0),
    }


def run_reference_arm(args, wl):
    """--impl reference: rank 0 only; each step = one encode + decode of a bounded sample."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    method, f, gen, n, desc = wl
    per_step = min(n, 20_000_000 if method != M_INT or gen[0] == "geom" else 5_000_000)
    sample = generate_host(gen, per_step, seed=0)
    codec, kind = cpu_codec()
    block = PA_MODEL_SYMS.get(args.workload, 0)
    for _ in range(args.warmup):
        codec.time(method, f, sample[: per_step // 8], 1, block)
    t0 = time.time()
    enc = dec = 0.0
    nbytes = 0
    for _ in range(args.steps):
        e, d, nbytes = codec.time(method, f, sample, 1, block)
        enc += e
        dec += d
    wall = time.time() - t0
    gbs = 8.0 * per_step * args.steps / (enc + dec)
    line = {
        "impl": "reference", "metric": METRIC, "value": gbs, "unit": UNIT, "n_gpus": args.gpus, "steps": args.steps,
        "warmup": args.warmup, "ms_per_step": (enc + dec) / args.steps / 1e6, "higher_is_better": True,
        "scaling": "strong" if args.workload in STRONG else "weak", "vs_baseline": None, "dtype": "u64", "data": "synthetic",
        "config": {"workload": desc, "method": method_label(method, f), "symbols_per_step": per_step,
                   "note": "bounded sample of the workload per step; CPU throughput does not depend on n"},
        "cpu_baseline": {"value": gbs, "unit": UNIT, "cores": 1, "kind": kind,
                         "sample": f"{per_step} symbols per step x {args.steps} steps, whole-stream encode + decode",
                         "encode_gbs": 4.0 * per_step * args.steps / enc, "decode_gbs": 4.0 * per_step * args.steps / dec,
                         "bits_per_symbol": 8.0 * nbytes / per_step, "host_cores_total": os.cpu_count()},
        "e2e": {"value": gbs, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0, "wall_s": wall,
    }
    emit(line)


# ------------------------------------------------------------------------------------------------
# our arm
# ------------------------------------------------------------------------------------------------
def blob_capacity(codec, method, n, pa=0):
    """Destination capacity: the library's worst-case bound for small inputs; 3 bytes per symbol (ANS: 4) for the
    2^30 / 2^32-symbol jobs, whose worst case (7 bytes per symbol) would not fit next to the data (the call fails with
    B200ANS_E_CAPACITY if the data turned out less compressible, it never overruns)."""
    bound = codec.pa_bound(method, pa, n) if pa else codec.bound(method, n)
    if n >= 1 << 28:
        bound = min(bound, (4 if method == M_INT else 3) * n + (64 << 20))
    return (bound + 15) // 16 * 16


def alg_bytes(n, c):
    """algorithmic bytes per launch; c = payload bytes per symbol"""
    return {"histogram": 4.0 * n, "encode": (4.0 + c) * n, "compact": 2.0 * c * n, "decode": (4.0 + c) * n}


def block_parity(codec, method, f, d_in, n, d_blob, nbytes, group, world, k=64, seed=0):
    """Outside the timed region: k random blocks (+ the first and the last) of this rank's container against the
    REFERENCE coder (oracle/_ref, unmodified headers) run on those blocks with the same global model -- byte for byte
    -- plus normalised frequencies and prelude.  The reference model is built from the global histogram (counted here
    with torch.bincount, not with the library) by the reference's own adjust_freqs.  ANSrfold (no histogram
    constructor in the shim): the reference DECODER, loaded from our prelude, must return the block's symbols."""
    import torch
    import torch.distributed as dist

    import _container
    from _libs import have_reference, reference

    if not have_reference():
        return {"checked": 0, "note": "oracle/_ref/libansref.so not present"}
    ref = reference()
    dev = d_in.device
    front = d_blob[: min(nbytes, 64 + (64 << 20))].cpu().numpy()
    hdr = codec_container_info(front)
    bs, nb, pre = hdr["block_syms"], hdr["n_blocks"], hdr["prelude_bytes"]
    table_at = 64 + (pre + 3) // 4 * 4
    payload_at = (table_at + 4 * nb + 15) // 16 * 16
    if front.size < payload_at:
        front = d_blob[:payload_at].cpu().numpy()
    sizes = front[table_at : table_at + 4 * nb].view(np.uint32).astype(np.int64)
    offs = np.concatenate([[0], np.cumsum((sizes + 15) // 16 * 16)[:-1]])
    prelude = front[64 : 64 + pre]
    out = {"blocks": int(nb), "block_syms": int(bs)}
    model = None
    if method != M_RFOLD:
        # the reference's counts: its own mapping (the host restatement is pinned against it), counted by torch
        if method == M_INT:
            mx = d_in.max().to(torch.int64).reshape(1)
            if world > 1:
                dist.all_reduce(mx, op=dist.ReduceOp.MAX, group=group)
            nbins = int(mx.item()) + 1
            sym = d_in
        else:
            nbins = 1280 if method == M_MSB else 1 << (f + 9)
            sym = None
        bins = torch.zeros(nbins, dtype=torch.int64, device=dev)
        chunk = 1 << 27
        for lo in range(0, n, chunk):
            v = d_in[lo : lo + chunk].to(torch.int64)
            if method != M_INT:
                v = fold_symbols_torch(v, method, f)
            bins += torch.bincount(v, minlength=nbins)[:nbins]
        if world > 1:
            dist.all_reduce(bins, op=dist.ReduceOp.SUM, group=group)
        hbins = bins.cpu().numpy().astype(np.uint64)
        largest = int(np.nonzero(hbins)[0][-1])
        model = ref.model_from_hist(method, f, hbins[: largest + 1] if method == M_INT else hbins)
        frame, nfreqs = model.info()
        ours_frame, ours_nf = codec.model_nfreqs()
        out["nfreqs_equal"] = bool(ours_frame == frame and np.array_equal(ours_nf, nfreqs))
        out["prelude_equal"] = bool(_container.prelude_equal(prelude, model.serialize()))
        out["frame_size"] = int(frame)
    dec = ref.model_load(method, f, prelude) if method == M_RFOLD else None
    rng = np.random.default_rng(seed)
    pick = sorted(set([0, nb - 1] + [int(x) for x in rng.integers(0, nb, k)]))
    bad = []
    for b in pick:
        lo, hi = b * bs, min(n, (b + 1) * bs)
        syms = d_in[lo:hi].cpu().numpy().view(np.uint32)
        pay = d_blob[payload_at + int(offs[b]) : payload_at + int(offs[b]) + int(sizes[b])].cpu().numpy()
        if model is not None:
            want = model.encode_block(syms)
            same = want.size == pay.size and np.array_equal(want, pay)
        else:
            same = np.array_equal(dec.decode_block(pay, hi - lo), syms)
        if not same:
            bad.append(b)
    out.update({"checked": len(pick), "mismatched": bad,
                "how": ("payload bytes == reference encode_symbol/flush_state loop on the block (ans_fold.hpp:250-278), model from the global histogram"
                        if model is not None else "reference decoder (model loaded from our prelude) returns the block's symbols"),
                "ok": not bad and out.get("nfreqs_equal", True) and out.get("prelude_equal", True)})
    return out


def codec_container_info(front):
    import ans_large_alphabet_b200 as P

    return P.host_api().container_info(front)


def fold_symbols_torch(v, method, f):
    """ans_fold_mapping<f> (ans_fold.hpp:38-50) / ans_msb_mapping (ans_msb.hpp:41-50) on an int64 tensor"""
    import torch

    if method == M_MSB:
        k = (v > 256).to(torch.int64) + (v > 65536).to(torch.int64) + (v > (1 << 24)).to(torch.int64)
        return (v >> (8 * k)) + 256 * k
    thres, step = 1 << (f + 7), 255 << (f - 1)
    k = torch.zeros_like(v)
    x = v.clone()
    for _ in range(3):
        big = x >= thres
        x = torch.where(big, x >> 8, x)
        k += big.to(torch.int64)
    return x + k * step


def north_star_summary(P, dev, stream, peak, names=("fold1", "fold2", "ans30"), steps=3):
    """Device-resident kernel times of the north star's 2^30-symbol targets (ANSfold-1 / ANSfold-2 on Zipf, ANS on
    geometric(0.01)), block size from b200ans_suggest_block_syms in [8192, 16384]: per kernel ms and fraction of the
    measured HBM peak, plus the byte-parity of sampled blocks against the reference."""
    import torch

    out = {}
    for name in names:
        method, f, gen, n, desc = WORKLOADS[name]
        codec = P.Codec(device=dev.index, block_syms=8192, stream=stream.cuda_stream)
        block_syms = codec.suggest_block_syms(n, 8192, 16384)
        codec.set_block_syms(block_syms)
        try:
            d_in = generate_device(gen, n, dev, seed=0)
            d_blob = torch.empty(blob_capacity(codec, method, n), dtype=torch.uint8, device=dev)
            d_out = torch.empty(n, dtype=torch.int32, device=dev)
            torch.cuda.synchronize(dev)
            acc = {"histogram": 0.0, "encode": 0.0, "compact": 0.0, "decode": 0.0}
            e0, e1, e2 = (torch.cuda.Event(enable_timing=True) for _ in range(3))
            enc_ms = dec_ms = 0.0
            for it in range(2 + steps):
                with torch.cuda.stream(stream):
                    e0.record(stream)
                    nb = codec.compress_device(method, f, d_blob.data_ptr(), d_blob.numel(), d_in.data_ptr(), n)
                    s1 = codec.stats()
                    e1.record(stream)
                    codec.decompress_device(method, f, d_out.data_ptr(), n, d_blob.data_ptr(), nb)
                    s2 = codec.stats()
                    e2.record(stream)
                stream.synchronize()
                if it >= 2:
                    acc["histogram"] += s1.ms_histogram / steps
                    acc["encode"] += s1.ms_encode / steps
                    acc["compact"] += s1.ms_compact / steps
                    acc["decode"] += s2.ms_decode / steps
                    enc_ms += e0.elapsed_time(e1) / steps
                    dec_ms += e1.elapsed_time(e2) / steps
            c = s1.payload_bytes / n
            alg = alg_bytes(n, c)
            names_k = {"histogram": "histogram_kernel", "encode": s1.encode_kernel.decode(), "compact": "compact_kernel",
                       "decode": s2.decode_kernel.decode()}
            with torch.cuda.stream(stream):
                par = block_parity(codec, method, f, d_in, n, d_blob, nb, None, 1, k=64)
            out[name] = {
                "workload": desc, "block_syms": block_syms, "roundtrip_ok": bool(torch.equal(d_in, d_out)), "bits_per_symbol": 8.0 * nb / n,
                "encode_ms": enc_ms, "decode_ms": dec_ms, "encode_gbs": 4.0 * n / (enc_ms * 1e6), "decode_gbs": 4.0 * n / (dec_ms * 1e6),
                "encode_pass_frac": (8.0 + c) * n / (enc_ms * 1e6) / peak, "decode_pass_frac": (4.0 + c) * n / (dec_ms * 1e6) / peak,
                "kernels": {k: {"kernel": names_k[k], "ms": v, "achieved_gbs": alg[k] / (v * 1e6), "frac": alg[k] / (v * 1e6) / peak}
                            for k, v in acc.items() if v > 0},
                "parity": par,
            }
            del d_in, d_blob, d_out
        except Exception as e:  # never let the extra block take the headline line down
            out[name] = {"error": repr(e)}
        finally:
            codec.close()
        torch.cuda.empty_cache()
    return out


def copy_only(stream, dev, h_in, d_in, h_blob, d_blob, h_out, d_out, nbh, reps, serial_encode=False, barrier=None):
    """The same bytes as one e2e step and nothing else: 4n in + container out (encode side), container in + 4n out
    (decode side), page-locked buffers, the two directions of each side on their own streams (PCIe is full duplex).
    Returns ms per step -- the floor any host-buffer implementation on this box has. `serial_encode`: the encode side's
    two copies one after the other, as the data dependence of a semi-static coder forces them (no byte of the container
    exists before the whole input has been counted)."""
    import torch

    side = torch.cuda.Stream(device=dev)
    t0, t1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    torch.cuda.synchronize(dev)
    for rep in range(reps + 1):  # (one untimed pass first; the ranks start the timed ones together)
        if rep == 1 or reps == 0:
            torch.cuda.synchronize(dev)
            if barrier is not None:
                barrier()
            t0.record(stream)
        side.wait_stream(stream)
        with torch.cuda.stream(stream):
            d_in.copy_(h_in, non_blocking=True)
        if serial_encode:
            side.wait_stream(stream)
        with torch.cuda.stream(side):
            h_blob[:nbh].copy_(d_blob[:nbh], non_blocking=True)
        stream.wait_stream(side)
        side.wait_stream(stream)
        with torch.cuda.stream(stream):
            d_blob[:nbh].copy_(h_blob[:nbh], non_blocking=True)
        with torch.cuda.stream(side):
            h_out.copy_(d_out, non_blocking=True)
        stream.wait_stream(side)
    t1.record(stream)
    torch.cuda.synchronize(dev)
    return t0.elapsed_time(t1) / reps


def run_ours(args, wl):
    import torch
    import torch.distributed as dist

    import ans_large_alphabet_b200 as P
    from ans_large_alphabet_b200 import sharded

    method, f, gen, n, desc = wl
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device -- b200ans has no CPU fallback")
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    group = None
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)
        group = dist.group.WORLD

    if args.symbols:
        n = args.symbols
    strong = args.workload in STRONG and not args.symbols
    n_job = n * world if not strong else n
    if strong:
        # one fixed job, contiguous shards of whole blocks' worth of symbols; 2^32 / N is exact for N = 1, 2, 4, 8
        n = n_job // world
    stream = torch.cuda.Stream(device=dev)
    codec = P.Codec(device=local, block_syms=args.block_syms or 8192, stream=stream.cuda_stream)
    if not args.block_syms:
        # the container's own parameter: blocks that fill whole rounds of the lane-per-block coders on this GPU
        # (from 6144 up: a 2^29-symbol shard -- the 2^32-symbol job on eight GPUs -- is exactly one round of 7168-symbol
        # blocks, where 8192 would leave a seventh of the warps without work)
        args.block_syms = codec.suggest_block_syms(n, 6144, 16384)
        codec.set_block_syms(args.block_syms)
        block_rule = "b200ans_suggest_block_syms(n, 6144, 16384)"
    else:
        block_rule = "fixed"
    pa = PA_MODEL_SYMS.get(args.workload, 0)
    if world > 1 and not pa:
        # the library's own NCCL communicator (ncclCommInitRank inside b200ans_comm_init; torch.distributed only carries
        # the unique id): the histogram all-reduce then runs inside b200ans_compress_sharded[_device], on the codec's stream
        sharded.init_comm(codec, group)
    d_in = generate_device(gen, n, dev, seed=rank)
    d_blob = torch.empty(blob_capacity(codec, method, n, pa), dtype=torch.uint8, device=dev)
    d_out = torch.empty(n, dtype=torch.int32, device=dev)
    torch.cuda.synchronize(dev)

    def sync():
        # the codec runs on `stream` and so do the torch ops around it (every step is inside torch.cuda.stream(stream)):
        # stream order is all the ordering the sharded path needs
        pass

    def barrier():
        if world > 1:
            dist.barrier(group)
        torch.cuda.synchronize(dev)

    kern = {"histogram": 0.0, "encode": 0.0, "compact": 0.0, "decode": 0.0, "model_host": 0.0, "allreduce": 0.0}

    def enc_dev():
        if pa:  # per-block models: nothing is shared between ranks
            return codec.pa_compress_device(method, f, pa, d_blob.data_ptr(), d_blob.numel(), d_in.data_ptr(), n)
        if world == 1:
            return codec.compress_device(method, f, d_blob.data_ptr(), d_blob.numel(), d_in.data_ptr(), n)
        return codec.compress_sharded_device(method, f, d_blob.data_ptr(), d_blob.numel(), d_in.data_ptr(), n)

    def dec_dev(nbytes):
        if pa:
            codec.pa_decompress_device(method, f, d_out.data_ptr(), n, d_blob.data_ptr(), nbytes)
        else:
            codec.decompress_device(method, f, d_out.data_ptr(), n, d_blob.data_ptr(), nbytes)

    def device_step(acc):
        with torch.cuda.stream(stream):
            nb = enc_dev()
            s = codec.stats()
            enc_model = s.ms_model
            dec_dev(nb)
            s2 = codec.stats()
        if acc is not None:
            acc["histogram"] += s.ms_histogram
            acc["encode"] += s.ms_encode
            acc["compact"] += s.ms_compact
            acc["decode"] += s2.ms_decode
            acc["model_host"] += enc_model + s2.ms_model
            acc["allreduce"] += s.ms_allreduce
        return nb, s, s2

    # ---- device-resident timing -------------------------------------------------------------
    for _ in range(args.warmup):
        device_step(None)
    sampler = ClockSampler(local)
    if rank == 0:
        sampler.start()
        time.sleep(0.25)
    launches0 = codec.stats().launches
    ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    barrier()
    t_wall0 = time.time()
    with torch.cuda.stream(stream):
        ev0.record(stream)
    nb = 0
    for _ in range(args.steps):
        nb, st1, st2 = device_step(kern)
    with torch.cuda.stream(stream):
        ev1.record(stream)
    barrier()
    t_wall1 = time.time()
    ms_total = ev0.elapsed_time(ev1)
    launches = codec.stats().launches - launches0
    # A timed region shorter than a few nvidia-smi periods: every rank keeps the same step running (untimed) for about
    # 1.5 s so that the sampler sees the clocks under this load.  The count comes from the max-over-ranks step time, so
    # all ranks agree on it (the sharded step holds a collective).
    tmax = torch.tensor([ms_total], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(tmax, op=dist.ReduceOp.MAX, group=group)
    if float(tmax.item()) < 1500.0:
        extra = int(min(400, math.ceil(1500.0 / max(float(tmax.item()) / args.steps, 0.05))))
        for _ in range(extra):
            device_step(None)
        torch.cuda.synchronize(dev)
        t_wall1 = time.time()
    clocks = sampler.stop(t_wall0, t_wall1) if rank == 0 else None
    ok = bool(torch.equal(d_in, d_out))
    st = codec.stats()
    payload_bytes, total_bytes, n_blocks = st1.payload_bytes, nb, st1.n_blocks
    frame, nsyms, prelude_bytes = st1.frame_size, st1.n_symbols_model, st1.prelude_bytes
    enc_name, dec_name = st1.encode_kernel.decode(), st2.decode_kernel.decode()

    # encode-only / decode-only split (events around each pass)
    enc_ms = dec_ms = 0.0
    e0, e1, e2 = (torch.cuda.Event(enable_timing=True) for _ in range(3))
    reps = max(1, min(args.steps, 5))
    for _ in range(reps):
        barrier()
        with torch.cuda.stream(stream):
            e0.record(stream)
            nb2 = enc_dev()
            e1.record(stream)
            dec_dev(nb2)
            e2.record(stream)
        stream.synchronize()
        enc_ms += e0.elapsed_time(e1)
        dec_ms += e1.elapsed_time(e2)
    enc_ms /= reps
    dec_ms /= reps

    # ---- byte parity of this step's container against the reference (outside every timed region) ----
    parity = None
    if not pa and not args.no_parity:
        with torch.cuda.stream(stream):
            if world > 1:  # the staged path leaves the global model in the codec; re-encode so the check sees this step's bytes
                nb = enc_dev()
            parity = block_parity(codec, method, f, d_in, n, d_blob, nb, group, world, k=64, seed=rank)
        torch.cuda.synchronize(dev)

    # ---- end to end: pinned host buffers through the host entry points ------------------------
    h_in = torch.empty(n, dtype=torch.int32).pin_memory()
    h_in.copy_(d_in)
    h_blob = torch.empty(d_blob.numel(), dtype=torch.uint8).pin_memory()
    h_out = torch.empty(n, dtype=torch.int32).pin_memory()
    torch.cuda.synchronize(dev)

    def e2e_step():
        if pa:
            nbh = codec.pa_compress_ptr(method, f, pa, h_blob.data_ptr(), h_blob.numel(), h_in.data_ptr(), n)
            codec.pa_decompress_ptr(method, f, h_out.data_ptr(), n, h_blob.data_ptr(), nbh)
            return nbh
        if world == 1:
            nbh = codec.compress_ptr(method, f, h_blob.data_ptr(), h_blob.numel(), h_in.data_ptr(), n)
            codec.decompress_ptr(method, f, h_out.data_ptr(), n, h_blob.data_ptr(), nbh)
            return nbh
        # host buffers in, host buffers out: the collective host-buffer entry point, then (no collective) the same
        # pipelined decode call as at N = 1
        nbh = codec.compress_sharded_ptr(method, f, h_blob.data_ptr(), h_blob.numel(), h_in.data_ptr(), n)
        codec.decompress_ptr(method, f, h_out.data_ptr(), n, h_blob.data_ptr(), nbh)
        return nbh

    for _ in range(min(args.warmup, 2)):
        e2e_step()
    e2e_steps = max(1, min(args.steps, 5))
    f0, f1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    barrier()
    tw0 = time.perf_counter()
    with torch.cuda.stream(stream):
        f0.record(stream)
    for _ in range(e2e_steps):
        nbh = e2e_step()
    with torch.cuda.stream(stream):
        f1.record(stream)
    barrier()
    e2e_wall_ms = (time.perf_counter() - tw0) * 1e3  # host wall clock, barrier to barrier: what a caller waits for
    e2e_dev_ms = f0.elapsed_time(f1)
    e2e_ms = max(e2e_wall_ms, e2e_dev_ms)
    e2e_ok = bool(torch.equal(h_in, h_out))
    # the same calls on PAGEABLE host memory (what the C++ adaptors get from std::vector, methods.hpp:535-546): N = 1, one
    # warm + one timed step, outside every other timed region
    pageable_ms, n_pg = 0.0, min(n, 1 << 28)  # (a bounded sample: at most 1 GiB of symbols in pageable memory)
    if world == 1 and not pa:
        p_in = torch.empty(n_pg, dtype=torch.int32)
        p_in.copy_(h_in[:n_pg])
        p_blob = torch.empty(blob_capacity(codec, method, n_pg, 0), dtype=torch.uint8)
        p_out = torch.empty(n_pg, dtype=torch.int32)
        for it in range(2):
            tp0 = time.perf_counter()
            nbp = codec.compress_ptr(method, f, p_blob.data_ptr(), p_blob.numel(), p_in.data_ptr(), n_pg)
            codec.decompress_ptr(method, f, p_out.data_ptr(), n_pg, p_blob.data_ptr(), nbp)
            pageable_ms = (time.perf_counter() - tp0) * 1e3
        e2e_ok = e2e_ok and bool(torch.equal(p_in, p_out))
        del p_in, p_blob, p_out
    d_out.zero_()
    barrier()
    copy_ms = copy_only(stream, dev, h_in, d_in, h_blob, d_blob, h_out, d_out, nbh, reps=min(3, e2e_steps), barrier=barrier)
    copy_serial_ms = copy_only(stream, dev, h_in, d_in, h_blob, d_blob, h_out, d_out, nbh, reps=min(3, e2e_steps), serial_encode=True, barrier=barrier)

    # ---- reduce over ranks: MAX time ----------------------------------------------------------
    t = torch.tensor([ms_total, e2e_ms, enc_ms, dec_ms, copy_ms, e2e_dev_ms, copy_serial_ms], dtype=torch.float64, device=dev)
    par_ok = True if parity is None else bool(parity.get("ok", True))
    okt = torch.tensor([int(ok and e2e_ok), int(par_ok)], dtype=torch.int32, device=dev)
    KEYS = ("histogram", "encode", "compact", "decode", "model_host", "allreduce")
    kt = torch.tensor([kern[k] for k in KEYS], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX, group=group)
        dist.all_reduce(okt, op=dist.ReduceOp.MIN, group=group)
        dist.all_reduce(kt, op=dist.ReduceOp.MAX, group=group)
    ms_total, e2e_ms, enc_ms, dec_ms, copy_ms, e2e_dev_ms, copy_serial_ms = (float(x) for x in t.tolist())
    all_ok, par_all_ok = bool(okt[0].item()), bool(okt[1].item())
    for k, v in zip(KEYS, kt.tolist()):
        kern[k] = float(v)

    if rank == 0:
        ms_step = ms_total / args.steps
        gb_step = 8.0 * n * world / 1e9
        value = gb_step / (ms_step / 1e3)
        c = payload_bytes / n
        peak, peak_src = measured_peak()
        per = {k: v / args.steps for k, v in kern.items()}
        alg = alg_bytes(n, c)
        knames = {"histogram": "histogram_kernel", "encode": enc_name, "compact": "compact_kernel", "decode": dec_name}
        kernels = {}
        for k in ("histogram", "encode", "compact", "decode"):
            if per[k] > 0:
                a = alg[k] / (per[k] * 1e6)
                kernels[k] = {"kernel": knames[k], "ms": per[k], "algorithmic_bytes": alg[k], "achieved_gbs": a, "frac": a / peak}
        dom = max(kernels, key=lambda k: per[k])
        roofline = {"bound": "hbm", "kernel": knames[dom], "achieved": kernels[dom]["achieved_gbs"], "peak": peak,
                    "unit": "GB/s", "frac": kernels[dom]["frac"], "traffic": None, "peak_source": peak_src,
                    "algorithmic_bytes_per_symbol": alg[dom] / n, "ms_per_launch": per[dom],
                    "note": "per launch on one GPU (max over ranks); algorithmic bytes = SURVEY 8(d) x the rank's symbols"}
        # DRAM bytes of that kernel from the committed `ncu --set full` capture of this workload (tools/ncu_traffic.py)
        try:
            with open(os.path.join(ROOT, "profiles", "traffic.json")) as fh:
                tr = json.load(fh).get(args.workload, {}).get(dom)
            if tr and not args.symbols and world == 1:
                roofline["traffic"] = tr["dram_bytes_per_launch"]
                roofline["traffic_source"] = f"ncu dram__bytes_read.sum + dram__bytes_write.sum, {tr['kernel']}, profiles/ ({tr['source']})"
                roofline["traffic_over_algorithmic"] = tr["dram_bytes_per_launch"] / alg[dom]
        except (OSError, ValueError, KeyError):
            pass
        # CPU baseline on a bounded sample of the same workload (rank 0, N = 1 only)
        cpu = None
        if world == 1 and not args.no_cpu:
            ns = min(n, args.cpu_sample)
            sample = d_in[:ns].cpu().numpy().view(np.uint32)
            cpu = cpu_time_sample(method, f, sample, runs=2, block=pa)
            sb = codec.pa_compress(method, f, pa, sample) if pa else codec.compress(method, f, sample)
            bps_sample = 8.0 * sb.size / ns
            cpu["ours_bits_per_symbol_same_sample"] = bps_sample
            cpu["size_overhead_vs_reference_pct"] = 100.0 * (bps_sample / cpu["bits_per_symbol"] - 1.0)
        h2d = 4 * n + nbh
        d2h = nbh + 4 * n
        line = {
            "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
            "ms_per_step": ms_step, "higher_is_better": True, "scaling": "strong" if strong else "weak", "vs_baseline": None,
            "dtype": "u64", "data": "synthetic",
            "config": {"workload": desc, "method": method_label(method, f), "symbols_job": n_job, "symbols_per_gpu": n,
                       "block_syms": pa or args.block_syms, "block_syms_rule": "one model per block" if pa else block_rule, "n_blocks_per_gpu": n_blocks,
                       "l2": "inputs (%.0f MB per GPU) larger than the 126 MB L2; no flush" % (4.0 * n / 1e6),
                       "model": ("pseudo-adaptive: one model per block, normalised on the device, no collective" if pa else
                                 "semi-static, global histogram" + (" via one ncclAllReduce inside b200ans_compress_sharded" if world > 1 else "")),
                       "frame_size": frame, "model_symbols": nsyms,
                       "frame_note": ("jobs of 2^31 symbols and more: the reference's 32-bit accumulators in cross_entropy (util.hpp:288-289) "
                                      "change adjust_freqs' stopping rule; the frame is the reference's for this histogram (SURVEY 3.8)"
                                      if n_job >= 1 << 31 else None)},
            "encode_gbs": 4.0 * n * world / (enc_ms * 1e6), "decode_gbs": 4.0 * n * world / (dec_ms * 1e6),
            "encode_ms": enc_ms, "decode_ms": dec_ms,
            "bits_per_symbol": 8.0 * total_bytes / n, "payload_bytes_per_symbol": c, "prelude_bytes": prelude_bytes,
            "roundtrip_ok": all_ok, "parity": dict(parity or {}, all_ranks_ok=par_all_ok),
            "roofline": roofline, "kernels": kernels, "host_model_ms_per_step": per["model_host"],
            "allreduce_ms_per_step": per["allreduce"],
            "cpu_baseline": cpu,
            "e2e": {"value": gb_step / (e2e_ms / e2e_steps / 1e3), "unit": UNIT, "h2d_bytes_per_step": h2d,
                    "d2h_bytes_per_step": d2h, "steps": e2e_steps, "ms_per_step": e2e_ms / e2e_steps,
                    "clock": "host wall clock, barrier to barrier, max over ranks", "device_ms_per_step": e2e_dev_ms / e2e_steps,
                    "pageable_gbs": (8.0 * n_pg / 1e9 / (pageable_ms / 1e3)) if pageable_ms > 0 else None,
                    "pageable_note": "the same two calls on pageable host memory (what the C++ adaptors hand over from std::vector), "
                                     f"one step over {n_pg} symbols of the workload",
                    "copy_only_ms": copy_ms, "copy_only_gbs": gb_step / (copy_ms / 1e3),
                    "copy_floor_ms": copy_serial_ms, "copy_floor_gbs": gb_step / (copy_serial_ms / 1e3),
                    "copy_floor_note": "as copy_only, but the encode side's copy-out after its copy-in: a semi-static coder cannot write a byte "
                                       "before it has counted the whole input, so this is the floor of b200ans_compress + b200ans_decompress",
                    "copy_only_note": "the same bytes per step with cudaMemcpyAsync only, both directions overlapped, all ranks at once",
                    "api": "b200ans_pa_compress + b200ans_pa_decompress (host buffers, pinned)" if pa
                    else "b200ans_compress + b200ans_decompress (host buffers, pinned)" if world == 1
                    else "b200ans_compress_sharded + b200ans_decompress per rank (host buffers, pinned)"},
            "gpu_launches": int(launches), "clocks": clocks,
        }
        if pa and world == 1:
            line["driver_remap"] = driver_remap_summary(codec, d_in, d_out, n, pa, stream)
        if world == 1 and not args.no_targets and not args.symbols:
            del d_in, d_blob, d_out, h_in, h_blob, h_out
            torch.cuda.empty_cache()
            line["north_star_2p30"] = north_star_summary(P, dev, stream, peak)
        emit(line)
    codec.close()
    if world > 1:
        dist.destroy_process_group()
    if not all_ok:
        raise SystemExit("bench.py: round trip FAILED")
    if not par_all_ok:
        raise SystemExit("bench.py: block parity against the reference FAILED")


def driver_remap_summary(codec, d_in, d_out, n, model_syms, stream):
    """Config 4 as the reference's DRIVER codes it (src/pseudo_adaptive.cpp:83-125): every block remapped to dense ranks,
    the block alphabet in front, ANSmsb of the ranks (b200ans_pa_remap_*; outside the timed region).  The driver's BPI
    counts the coded ranks only (:122-125 leave the alphabets of full blocks out): ours = the inner container; the
    driver's own on the first 64 blocks through the unmodified reference."""
    import torch

    from _libs import have_reference, reference

    try:
        cap = codec.pa_bound(M_MSB, model_syms, n) + n * 4 + (n // model_syms + 2) * 48 + 256
        d_blob = torch.empty((cap + 15) // 16 * 16, dtype=torch.uint8, device=d_in.device)
        best = None
        for _ in range(3):
            with torch.cuda.stream(stream):
                stream.synchronize()
                t0 = time.perf_counter()
                nb = codec.pa_remap_compress_device(M_MSB, 0, model_syms, d_blob.data_ptr(), d_blob.numel(), d_in.data_ptr(), n)
                stream.synchronize()
                t1 = time.perf_counter()
                codec.pa_remap_decompress_device(M_MSB, 0, d_out.data_ptr(), n, d_blob.data_ptr(), nb)
                stream.synchronize()
                t2 = time.perf_counter()
            if best is None or t2 - t0 < best[0]:
                best = (t2 - t0, t1 - t0, t2 - t1)
        inner = int(np.frombuffer(d_blob[32:40].cpu().numpy().tobytes(), dtype=np.uint64)[0])
        out = {"coder": "per-block dense-rank remap + ANSmsb (b200ans_pa_remap_compress_device)", "model_syms": model_syms,
               "roundtrip_ok": bool(torch.equal(d_in, d_out)), "encode_ms": best[1] * 1e3, "decode_ms": best[2] * 1e3,
               "bits_per_symbol_with_alphabets": 8.0 * nb / n, "bits_per_symbol_coded_ranks": 8.0 * inner / n}
        if have_reference():
            ref = reference()
            host = d_in[: 64 * model_syms].cpu().numpy().astype(np.uint32)
            ranks_bytes = alpha_bytes = 0
            for b in range(64):
                blob, ab, _ = ref.pa_remap_block(M_MSB, 0, host[b * model_syms : (b + 1) * model_syms])
                ranks_bytes += blob.size - ab
                alpha_bytes += ab
            out["reference_driver_bpi"] = 8.0 * ranks_bytes / (64 * model_syms)
            out["reference_bits_per_symbol_with_alphabets"] = 8.0 * (ranks_bytes + alpha_bytes) / (64 * model_syms)
            out["reference_sample"] = "the first 64 blocks through run<ANSmsb>'s loop (oracle/ref_shim.cpp ref_pa_remap_block)"
        return out
    except Exception as e:  # never lose the line over the extra
        return {"error": str(e)}


_REAL_STDOUT = None


def emit(line):
    """the ONE JSON line of the contract, on the process's real stdout"""
    text = json.dumps(line) + "\n"
    if _REAL_STDOUT is None:
        sys.stdout.write(text)
        sys.stdout.flush()
    else:
        os.write(_REAL_STDOUT, text.encode())


def main():
    # Libraries write to fd 1 behind Python's back (NCCL prints its version banner there at the first collective):
    # everything but the JSON line goes to stderr.
    global _REAL_STDOUT
    sys.stdout.flush()
    _REAL_STDOUT = os.dup(1)
    os.dup2(2, 1)
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--workload", default=DEFAULT_WORKLOAD, choices=sorted(WORKLOADS))
    ap.add_argument("--symbols", type=int, default=0, help="override the workload's symbol count (per GPU)")
    ap.add_argument("--block-syms", type=int, default=0, help="symbols per block; 0 = the workload's default")
    ap.add_argument("--no-targets", action="store_true", help="skip the 2^30-symbol north-star summary")
    ap.add_argument("--cpu-sample", type=int, default=100_000_000)
    ap.add_argument("--no-cpu", action="store_true")
    ap.add_argument("--no-parity", action="store_true", help="skip the byte comparison of sampled blocks with the reference")
    args = ap.parse_args()
    args.warmup = max(args.warmup, 3) if args.impl == "ours" else args.warmup
    wl = WORKLOADS[args.workload]
    if not args.block_syms:
        args.block_syms = BLOCK_SYMS.get(args.workload, 0)  # 0: chosen from the shard size once the codec exists
    if args.impl == "reference":
        run_reference_arm(args, wl)
    else:
        run_ours(args, wl)


if __name__ == "__main__":
    main()
