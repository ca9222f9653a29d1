"""bench.py -- the Serra09 all-pairs scoring path on N B200s, BASELINE.json's metric
("song pairs/sec and Qmax GCUPS ... vs host-CPU reference").

  python bench.py --gpus N --steps K --warmup W            (N > 1: launched under torchrun)
  python bench.py --impl reference --gpus N --steps K --warmup W

Workload (config.workload): BASELINE configs[1], the whatisacover-subset-shaped synthetic HPCP
set (10 000 songs, 5000 cliques x 2, seed 10000, lengths lognormal around 450 frames), m = 9,
kappa = 0.095.  A STEP is one pass of the hot path (OTI -> cross-similarity -> mutual-kNN CRP
-> Qmax + Dmax) over one batch of song pairs drawn uniformly (seeded) from the all-pairs list;
every step scores a different batch and the float64 song arena (~0.45 GB) is larger than L2.
At N > 1 each rank scores its own batch of the same size (weak scaling, no data-path
collective: pairs are independent); `value` = all ranks' pairs / max-over-ranks time.

`value` times the C-ABI call with the songs already resident in HBM; `e2e` times the plugin call
(Serra09.similarity) with the host->device copy of the song arena from pinned host memory, the
pair list and the device->host read of the scores inside the timed region.
"""
import os

os.environ.setdefault("OPENBLAS_NUM_THREADS", "1")
os.environ.setdefault("OMP_NUM_THREADS", "1")
os.environ.setdefault("MKL_NUM_THREADS", "1")

import argparse  # noqa: E402
import json  # noqa: E402
import subprocess  # noqa: E402
import sys  # noqa: E402
import time  # noqa: E402

import numpy as np  # noqa: E402

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

M_WIN, KAPPA = 9, 0.095
FP64_LANES_PER_SM = 64   # B200: FP64 SIMT at half the FP32 rate
N_SM = 148


def parse_args():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", choices=("b200", "reference"), default="b200")
    ap.add_argument("--songs", type=int, default=10000, help="songs of the whatisacover-shaped set")
    ap.add_argument("--pairs-per-step", type=int, default=200000, help="pairs per rank per step")
    ap.add_argument("--cpu-pairs", type=int, default=0,
                    help="pairs of the CPU sample (0: 800 per core in the default arm = 15-25 s of CPU work, "
                         "150 per core and step in the reference arm)")
    ap.add_argument("--sweep-gen", type=int, default=0, help="force a sweep generation (2 .. 6); 0 = library default")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    return ap.parse_args()


def make_workload(n_songs):
    from acoss_b200 import synth
    songs, labels = synth.make_song_set("whatisacover", seed=10000, n_songs=n_songs)
    return songs, labels


def step_pairs(n_songs, per_step, step, rank, world, seed=4242):
    """Batch of (i < j) pairs for (step, rank): uniform over the upper triangle, seeded."""
    rng = np.random.default_rng([seed, step, rank])
    i = rng.integers(0, n_songs, size=per_step + per_step // 8 + 16)
    j = rng.integers(0, n_songs, size=i.size)
    keep = i != j
    i, j = i[keep][:per_step], j[keep][:per_step]
    lo, hi = np.minimum(i, j), np.maximum(i, j)
    return np.stack([lo, hi], axis=1).astype(np.int32)


def pair_work(lengths, pairs, m=M_WIN):
    """(cells M*N, Qmax cell updates (M-2)(N-2), algorithmic bytes) summed over a pair batch."""
    M = np.maximum(lengths[pairs[:, 0]].astype(np.int64) - m + 1, 0)
    N = np.maximum(lengths[pairs[:, 1]].astype(np.int64) - m + 1, 0)
    cells = int((M * N).sum())
    qcells = int((np.maximum(M - 2, 0) * np.maximum(N - 2, 0)).sum())
    bins = 12
    abytes = int((4 * bins * (lengths[pairs[:, 0]].astype(np.int64) + lengths[pairs[:, 1]])).sum() + 8 * len(pairs))
    return cells, qcells, abytes


# ---------------------------------------------------------------------------- CPU arm
_CPU_SONGS = None


def _cpu_chunk(chunk):
    from oracle import crp_oracle, dp
    lib = dp.ref_lib() or dp.oracle_lib()
    out = []
    for i, j in chunk:
        Si, Sj = _CPU_SONGS[int(i)], _CPU_SONGS[int(j)]
        r = crp_oracle.serra09_pair(Si["chroma"].astype(np.float64), Sj["chroma"].astype(np.float64),
                                    Si["gchroma"], Sj["gchroma"], M_WIN, KAPPA, True, lib)
        out.append((r["qmax"], r["dmax"]))
    return out


def cpu_reference_run(songs, pairs, cores):
    """The reference's CPU path (CRPUtils restated in oracle/crp_oracle.py + its compiled
    SequenceAlignment.c when oracle/_ref travelled, else the C restatement) on all host cores,
    45 chunks like CoverAlgorithm.py:169.  Returns seconds."""
    import multiprocessing as mp
    global _CPU_SONGS
    _CPU_SONGS = songs
    chunks = [c for c in np.array_split(pairs, min(45, max(1, len(pairs)))) if len(c)]
    ctx = mp.get_context("fork")
    with ctx.Pool(cores) as pool:
        pool.map(_cpu_chunk, [chunks[0][:1]] * cores)  # warm the workers (imports, library load)
        t0 = time.perf_counter()
        pool.map(_cpu_chunk, chunks, chunksize=1)
        dt = time.perf_counter() - t0
    return dt


def cpu_sample_desc(n_pairs, cores):
    from oracle import dp
    return ("%d uniformly drawn pairs of the same workload; CRP half = numpy restatement of CRPUtils.py, "
            "DP half = %s; multiprocessing over %d processes, 45 chunks, OPENBLAS_NUM_THREADS=1"
            % (n_pairs, "the reference's SequenceAlignment.c compiled -Ofast (oracle/_ref)" if dp.ref_lib() is not None
               else "C restatement (oracle/seqalign_oracle.c)", cores))


# ---------------------------------------------------------------------------- clocks
class ClockSampler:
    Q = "clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap"

    def __init__(self, index):
        self.index = index
        self.proc = None

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", "-i", str(self.index), "--query-gpu=" + self.Q,
                                          "--format=csv,noheader,nounits", "-lms", "100"],
                                         stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
        except Exception:
            self.proc = None

    def stop(self):
        if self.proc is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        self.proc.terminate()
        try:
            out, _ = self.proc.communicate(timeout=5)
        except Exception:
            self.proc.kill()
            out = ""
        sm, mx, reasons = [], [], set()
        for line in out.strip().splitlines():
            f = [x.strip() for x in line.split(",")]
            if len(f) < 7:
                continue
            try:
                sm.append(float(f[0]))
                mx.append(float(f[1]))
            except ValueError:
                continue
            for name, v in zip(("hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"), f[3:7]):
                if v.lower().startswith("active"):
                    reasons.add(name)
        if not sm:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["no samples"]}
        return {"sm_mhz": float(np.median(sm)), "sm_max_mhz": float(max(mx)), "reasons": sorted(reasons), "samples": len(sm)}


# ---------------------------------------------------------------------------- reference arm
def run_reference(args, rank, world):
    """--impl reference: the reference's own CPU implementation of the path on this box's host
    cores (rank 0 only).  Each step is a bounded sample of the same workload."""
    if rank != 0:
        return
    songs, _ = make_workload(args.songs)
    lengths = np.array([s["chroma"].shape[1] for s in songs])
    cores = os.cpu_count() or 1
    n = args.cpu_pairs or max(64, 150 * cores)
    times, npairs, qcells = [], 0, 0
    for s in range(args.warmup + args.steps):
        pairs = step_pairs(args.songs, n, 10_000 + s, 0, 1)
        dt = cpu_reference_run(songs, pairs, cores)
        if s >= args.warmup:
            times.append(dt)
            npairs += len(pairs)
            qcells += pair_work(lengths, pairs)[1]
    total = sum(times)
    value = npairs / total
    line = {
        "impl": "reference", "metric": "song_pairs_per_sec", "value": value, "unit": "pairs/s",
        "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup, "ms_per_step": 1e3 * total / max(args.steps, 1),
        "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f64", "data": "synthetic",
        "config": workload_config(args, n),
        "qmax_gcups_whole_path": qcells / total / 1e9,
        "cpu_baseline": {"value": value, "unit": "pairs/s", "cores": cores, "kind": "port",
                         "sample": cpu_sample_desc(n, cores) + "; per step"},
        "e2e": {"value": value, "unit": "pairs/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
    }
    print(json.dumps(line), flush=True)


def workload_config(args, pairs_per_step):
    return {"workload": "whatisacover-subset-shaped synthetic HPCP (BASELINE configs[1]): %d songs, 12 bins, "
                        "lognormal lengths median 450 frames, all-pairs list sampled uniformly" % args.songs,
            "songs": args.songs, "pairs_per_step_per_gpu": pairs_per_step, "m": M_WIN, "kappa": KAPPA,
            "total_pairs_in_full_job": args.songs * (args.songs - 1) // 2,
            "l2_policy": "inputs larger than L2 (0.45 GB float64 song arena, a new pair batch every step)"}


# ---------------------------------------------------------------------------- B200 arm
def run_b200(args, rank, world, local):
    import torch
    import torch.distributed as dist
    from acoss_b200 import _native, parallel
    from acoss_b200.serra09 import Serra09

    distributed = world > 1

    def barrier():
        torch.cuda.synchronize()
        if distributed:
            dist.barrier()
        torch.cuda.synchronize()

    # songs: generated on rank 0, shared once over NCCL (the run's only feature collective)
    cpu_base = None
    if rank == 0:
        songs, labels = make_workload(args.songs)
        packed = _native.pack_songs(songs, pinned=False)
        if world == 1 and not args.no_cpu_baseline:
            # CPU arm first: its worker processes are forked before this process touches CUDA
            cores = os.cpu_count() or 1
            n = args.cpu_pairs or max(64, 800 * cores)
            secs = cpu_reference_run(songs, step_pairs(args.songs, n, 10_000, 0, 1), cores)
            cpu_base = {"value": n / secs, "unit": "pairs/s", "cores": cores, "kind": "port",
                        "sample": cpu_sample_desc(n, cores), "seconds": secs}
    else:
        songs, labels, packed = None, None, None
    if distributed:
        packed = parallel.broadcast_songs(packed, src=0)
        if rank != 0:
            frames, offs, lens, g = packed
            songs = [{"chroma": np.ascontiguousarray(frames[o:o + n].T), "gchroma": g[k]}
                     for k, (o, n) in enumerate(zip(offs, lens))]
            labels = ["clique_%05d" % k for k in range(len(songs))]
    if not torch.cuda.is_available():
        raise SystemExit("bench.py --impl b200 needs a CUDA device (there is no CPU fallback)")
    torch.cuda.set_device(local)
    lengths = np.asarray(packed[2])
    alg = Serra09(songs=songs, labels=labels, shortname="bench", do_memmaps=False, do_mfcc=False, device=local)
    alg.preload()
    ctx, slot = alg._context_for("chroma", np.arange(args.songs))
    if args.sweep_gen:
        ctx.set_option("sweep_gen", args.sweep_gen)
    B = args.pairs_per_step

    # ---- value: C-ABI call, songs resident in HBM
    for s in range(args.warmup):
        ctx.serra09_pairs(step_pairs(args.songs, B, s, rank, world), m=M_WIN, kappa=KAPPA)
    batches = [step_pairs(args.songs, B, args.warmup + s, rank, world) for s in range(args.steps)]
    sampler = ClockSampler(local)
    ctx.reset_counters()
    stage = {"oti": 0.0, "thresholds": 0.0, "crp": 0.0, "dp": 0.0, "total": 0.0}
    barrier()
    if rank == 0:
        sampler.start()
    t0 = time.perf_counter()
    for pairs in batches:
        ctx.serra09_pairs(pairs, m=M_WIN, kappa=KAPPA)   # synchronises its stream before returning
        for k, v in ctx.last_stage_ms().items():
            stage[k] += v
    barrier()
    dt = time.perf_counter() - t0
    clocks = sampler.stop() if rank == 0 else None
    launches = ctx.kernel_launches()

    # ---- e2e: plugin call, H2D of the song arena + pairs and D2H of the scores inside the region
    for s in range(2):
        alg.drop_device_arena()
        alg.similarity(step_pairs(args.songs, B, 500 + s, rank, world))
    e2e_batches = [step_pairs(args.songs, B, 600 + s, rank, world) for s in range(args.steps)]
    barrier()
    t1 = time.perf_counter()
    for pairs in e2e_batches:
        alg.drop_device_arena()
        sims = alg.similarity(pairs)
        if distributed:
            parallel.gather_scores(pairs, sims)   # score tiles back to rank 0
    barrier()
    dt_e2e = time.perf_counter() - t1

    tmax = torch.tensor([dt, dt_e2e], dtype=torch.float64, device="cuda")
    work = torch.tensor([float(x) for b in batches for x in pair_work(lengths, b)], dtype=torch.float64,
                        device="cuda").reshape(-1, 3).sum(0)
    if distributed:
        dist.all_reduce(tmax, op=dist.ReduceOp.MAX)
        dist.all_reduce(work, op=dist.ReduceOp.SUM)
    dt, dt_e2e = float(tmax[0]), float(tmax[1])
    cells, qcells, abytes = [float(x) for x in work]
    if rank != 0:
        return
    total_pairs = B * args.steps * world
    value = total_pairs / dt
    e2e_value = total_pairs / dt_e2e
    h2d = alg.host_arena_bytes() + B * 8
    d2h = B * 20
    # roofline of the dominant kernel group (k_sweep: thresholds x2 orientations + CRP), rank 0's launches
    peaks = {}
    try:
        peaks = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
    except Exception:
        pass
    my_cells, my_q, my_bytes = [float(sum(x)) for x in zip(*[pair_work(lengths, b) for b in batches])]
    sweep_s = (stage["thresholds"] + stage["crp"]) / 1e3
    sm_max = (clocks or {}).get("sm_max_mhz") or peaks.get("sm_max_mhz") or 1965.0
    fp64_peak = N_SM * FP64_LANES_PER_SM * 2 * sm_max * 1e6 / 1e12
    achieved_tf = 26.0 * my_cells / sweep_s / 1e12 if sweep_s > 0 else None
    hbm_peak = peaks.get("hbm_gbs", 6650.0)
    line = {
        "metric": "song_pairs_per_sec", "value": value, "unit": "pairs/s", "n_gpus": world, "steps": args.steps,
        "warmup": args.warmup, "ms_per_step": 1e3 * dt / args.steps, "higher_is_better": True, "scaling": "weak",
        "vs_baseline": None, "dtype": "f64", "data": "synthetic", "config": workload_config(args, B),
        "qmax_gcups_whole_path": qcells / dt / 1e9,
        "qmax_dmax_gcups_dp_kernel": (2.0 * my_q / (stage["dp"] / 1e3) / 1e9) if stage["dp"] > 0 else None,
        "cells_per_sec": cells / dt,
        "device_ms_per_step": {k: v / args.steps for k, v in stage.items()},
        "clocks": clocks,
        "e2e": {"value": e2e_value, "unit": "pairs/s", "h2d_bytes_per_step": int(h2d), "d2h_bytes_per_step": int(d2h),
                "ms_per_step": 1e3 * dt_e2e / args.steps},
        "gpu_launches": int(launches),
        "roofline": {"bound": "tensor", "achieved": achieved_tf, "peak": fp64_peak, "unit": "TFLOP/s",
                     "frac": (achieved_tf / fp64_peak) if achieved_tf else None, "traffic": None,
                     "kernel": "k_sweep6 (FP64 tensor-core sweep: separators of one orientation + CRP of the other); "
                               "k_sweep5 / k_sweep4 (FP64 SIMT) for column songs over 631 frames",
                     "note": "FP64 roofline, NOT the bf16 figure of MEASURED_PEAKS.json (which has no FP64 entry): "
                             "peak = 148 SM x 64 FP64 FMA/cycle x 2 x sm_max_mhz, the rate measured for both DMMA "
                             "(mma.sync.m8n8k4.f64, profiles/exp/ubench/dmma_rate.cu: 16 cycles per instruction per SM "
                             "sub-partition) and DFMA (profiles/exp/ubench/fp64_lat.cu); algorithmic 26 flop/cell counted "
                             "ONCE although every cell is computed twice (once per orientation); achieved = flops of all "
                             "sweep launches of the timed steps / their summed device time (CUDA events on the launching "
                             "stream); the sweeps are bound by the integer selection, not by the FP64 pipe (DESIGN.md 7)"},
        "roofline_hbm": {"bound": "hbm", "achieved": (my_bytes / sweep_s / 1e9) if sweep_s > 0 else None, "peak": hbm_peak,
                         "unit": "GB/s", "frac": (my_bytes / sweep_s / 1e9 / hbm_peak) if sweep_s > 0 else None,
                         "note": "compulsory bytes only (both songs' frames + scores); the path is compute-bound by design, "
                                 "a high number here would mean the CRP round-trips (peak: measured copy bandwidth)"},
    }
    if cpu_base is not None:
        line["cpu_baseline"] = cpu_base
    print(json.dumps(line), flush=True)
    alg.close()


def main():
    args = parse_args()
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if args.impl == "reference":
        run_reference(args, rank, world)
        return
    if world > 1:
        from acoss_b200 import parallel
        parallel.init_distributed()
    run_b200(args, rank, world, local)
    if world > 1:
        import torch.distributed as dist
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
