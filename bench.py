"""bench.py -- the Serra09 all-pairs scoring path on N B200s, BASELINE.json's metric
("song pairs/sec and Qmax GCUPS ... vs host-CPU reference").

  python bench.py --gpus N --steps K --warmup W            (N > 1: launched under torchrun)
  python bench.py --impl reference --gpus N --steps K --warmup W
  python bench.py --config benchmark ...                   (BASELINE configs[2]: 15 000 songs, cliques of 13)
  python bench.py --config benchmark --full-job            (the whole 112 492 500-pair matrix once; not a bench line)

Workload (config.workload): by default BASELINE configs[1], the whatisacover-subset-shaped synthetic HPCP set
(10 000 songs, 5000 cliques x 2, seed 10000, lengths lognormal around 450 frames), m = 9, kappa = 0.095.
The all-pairs matrix is cut into T x T song tiles (acoss_b200.parallel: upper triangle, T = 32 here).  A STEP is one
pass of the hot path (OTI -> cross-similarity -> mutual-kNN CRP -> Qmax + Dmax) over a fixed-size set of tiles
(--pairs-per-step pairs IN TOTAL, a different seeded tile set every step); the float64 song arena (~0.45 GB) is
larger than L2.  STRONG scaling: at N > 1 the SAME tile set is dealt to the N ranks by estimated cost
(longest-processing-time greedy), `value` = the step's pairs / max-over-ranks time.  No data-path collective:
pairs are independent.

`value` times the C-ABI call with the songs already resident in HBM.  `e2e` times the plugin-level distributed pass
with HOST buffers: every step each rank copies ITS shard of the songs from pinned host memory to its device, the
shards are all-gathered over NCCL (device to device), every rank scores its tiles, the float32 score tiles are
gathered on rank 0, read back and placed in the host N x N matrices.
"""
import os

os.environ.setdefault("OPENBLAS_NUM_THREADS", "1")
os.environ.setdefault("OMP_NUM_THREADS", "1")
os.environ.setdefault("MKL_NUM_THREADS", "1")

import argparse  # noqa: E402
import contextlib  # noqa: E402
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


CONFIGS = {
    # name: (synth shape, songs, seed, description)
    "covers80": ("covers80", 160, 80, "covers80-shaped synthetic HPCP (BASELINE configs[0]: 80 cliques of 2, the reference's own CPU-runnable case)"),
    "whatisacover": ("whatisacover", 10000, 10000, "whatisacover-subset-shaped synthetic HPCP (BASELINE configs[1])"),
    "benchmark": ("benchmark", 15000, 15000, "Da-TACOS benchmark-subset-shaped synthetic HPCP (BASELINE configs[2]: 1000 cliques x 13 + 2000 singletons)"),
    "stress": ("covers80", 96, 4096, "long-song stress (BASELINE configs[4]): 4096-frame HPCP songs plus a mixed 1024..4096 set, m = 20, "
                                     "variable-length batching"),
    "earlysnf": ("benchmark", 2000, 15000, "EarlySNF fused-matrix alignment over benchmark-subset shapes (BASELINE configs[3]): chroma + mfcc kernels "
                                           "fused by similarity network fusion, fused cross block -> mutual kNN -> Qmax / Dmax"),
}
TILE = 32  # song tile edge of the bench's pair tiles (1024 pairs per off-diagonal tile)


def parse_args():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", choices=("b200", "reference"), default="b200")
    ap.add_argument("--config", choices=sorted(CONFIGS), default="whatisacover")
    ap.add_argument("--songs", type=int, default=0, help="songs of the set (0: the config's own count)")
    ap.add_argument("--pairs-per-step", type=int, default=1200000, help="pairs per step, IN TOTAL over all GPUs (strong scaling)")
    ap.add_argument("--cpu-pairs", type=int, default=0,
                    help="pairs of the CPU sample (0: 800 per core in the default arm = 15-25 s of CPU work, "
                         "150 per core and step in the reference arm)")
    ap.add_argument("--sweep-gen", type=int, default=0, help="force a sweep generation (2, 4 or 7); 0 = library default")
    ap.add_argument("--all-pairs", action="store_true",
                    help="--impl reference: one step = every unique pair of the set (covers80: 12 720), not a sample")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--pair-order", choices=("tiles", "shuffled", "ascending"), default="tiles",
                    help="experiment: order of a rank's pair list inside the value region (tiles = as dealt, most expensive tiles first)")
    ap.add_argument("--opt", action="append", default=[], metavar="NAME=VALUE",
                    help="library tuning knob for A/B runs (acoss_set_option), e.g. --opt dp_packed=0")
    ap.add_argument("--mfcc", action="store_true", help="also time the default plugin call (chroma + mfcc halves) once")
    ap.add_argument("--full-job", action="store_true", help="score the whole all-pairs matrix once and report its wall time")
    args = ap.parse_args()
    if not args.songs:
        args.songs = CONFIGS[args.config][1]
    if args.config == "covers80" and args.pairs_per_step == 1200000:
        args.pairs_per_step = args.songs * (args.songs - 1) // 2
    if args.config == "stress":
        global M_WIN, TILE
        M_WIN, TILE = 20, 8          # BASELINE configs[4]: m = 20; 64-pair tiles (a 4096 x 4096 pair is 16.6 M cells)
        if args.pairs_per_step == 1200000:
            args.pairs_per_step = 2000
    return args


def make_workload(args):
    from acoss_b200 import synth
    shape, _, seed, _ = CONFIGS[args.config]
    if args.config == "stress":   # half of the songs at 4096 frames, half drawn from 1024..4096
        half = args.songs // 2
        songs, labels = synth.make_song_set(shape, seed=seed, n_songs=half, fixed_length=4096)
        s2, l2 = synth.make_song_set(shape, seed=seed + 1, n_songs=args.songs - half, length_range=(1024, 4096))
        return songs + s2, labels + ["m" + x for x in l2]
    songs, labels = synth.make_song_set(shape, seed=seed, n_songs=args.songs)
    return songs, labels


def step_pairs(n_songs, per_step, step, rank, world, seed=4242):
    """Batch of (i < j) pairs for (step, rank): uniform over the upper triangle, seeded (CPU arm, experiments)."""
    rng = np.random.default_rng([seed, step, rank])
    i = rng.integers(0, n_songs, size=per_step + per_step // 8 + 16)
    j = rng.integers(0, n_songs, size=i.size)
    keep = i != j
    i, j = i[keep][:per_step], j[keep][:per_step]
    lo, hi = np.minimum(i, j), np.maximum(i, j)
    return np.stack([lo, hi], axis=1).astype(np.int32)


class TileSets:
    """Seeded tile subsets of the all-pairs tile list: step s scores tiles(s), about `per_step` pairs in total."""

    def __init__(self, n_songs, lengths, per_step, T=None, seed=4242):
        from acoss_b200 import parallel
        T = TILE if T is None else T
        self.par, self.n, self.T, self.per_step, self.seed, self.lengths = parallel, n_songs, T, per_step, seed, lengths
        self.tiles = parallel.tile_list(n_songs, T)
        nt = (n_songs + T - 1) // T
        edge = np.minimum((np.arange(nt) + 1) * T, n_songs) - np.arange(nt) * T
        a = np.array([t[0] for t in self.tiles]); b = np.array([t[1] for t in self.tiles])
        self.count = np.where(a != b, edge[a] * edge[b], edge[a] * (edge[a] - 1) // 2)
        self.costs = parallel.tile_costs(lengths, T, self.tiles, M_WIN)

    def tile_ids(self, step):
        rng = np.random.default_rng([self.seed, step])
        perm = rng.permutation(len(self.tiles))
        csum = np.cumsum(self.count[perm])
        k = int(np.searchsorted(csum, self.per_step)) + 1
        return np.sort(perm[:k])

    def shares(self, step, world):
        """(tile ids, per-rank index arrays into them, per-rank estimated load)"""
        ids = self.tile_ids(step)
        sh, load = self.par.assign_tiles(self.costs[ids], world)
        return ids, sh, load

    def rank_pairs(self, step, world, rank):
        ids, sh, load = self.shares(step, world)
        pairs = self.par.tiles_pairs(self.n, self.T, [self.tiles[ids[t]] for t in sh[rank]])
        return pairs, int(self.count[ids].sum()), float(load.max() / max(load.mean(), 1e-30))


def pair_work(lengths, pairs, m=None):
    """(cells M*N, Qmax cell updates (M-2)(N-2), algorithmic bytes) summed over a pair batch."""
    m = M_WIN if m is None else m
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
    songs, _ = make_workload(args)
    lengths = np.array([s["chroma"].shape[1] for s in songs])
    cores = os.cpu_count() or 1
    n = args.cpu_pairs or max(64, 150 * cores)
    times, npairs, qcells = [], 0, 0
    if args.all_pairs:
        iu = np.triu_indices(args.songs, 1)
        every = np.stack(iu, axis=1).astype(np.int32)
        n = len(every)
    for s in range(args.warmup + args.steps):
        pairs = every if args.all_pairs else step_pairs(args.songs, n, 10_000 + s, 0, 1)
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
        "higher_is_better": True, "scaling": "strong", "vs_baseline": None, "dtype": "f64", "data": "synthetic",
        "config": workload_config(args),
        "cpu_pairs_per_step": n,
        "qmax_gcups_whole_path": qcells / total / 1e9,
        "cpu_baseline": {"value": value, "unit": "pairs/s", "cores": cores, "kind": "port",
                         "sample": (("all %d unique pairs of the set; " % n) + cpu_sample_desc(n, cores).split("; ", 1)[1]
                                    if args.all_pairs else cpu_sample_desc(n, cores)) + "; per step"},
        "e2e": {"value": value, "unit": "pairs/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
    }
    print(json.dumps(line), flush=True)


def workload_config(args):
    desc = CONFIGS[args.config][3]
    stress = args.config == "stress"
    lens = "half at 4096 frames, half uniform in 1024..4096" if stress else "lognormal lengths median 450 frames"
    return {"workload": "%s: %d songs, 12 bins, %s; a step = a seeded set of %d x %d song "
                        "tiles of the all-pairs matrix, about %d pairs in total, dealt to the GPUs by estimated cost"
                        % (desc, args.songs, lens, TILE, TILE, args.pairs_per_step),
            "songs": args.songs, "pairs_per_step_total": args.pairs_per_step, "tile": TILE, "m": M_WIN, "kappa": KAPPA,
            "total_pairs_in_full_job": args.songs * (args.songs - 1) // 2,
            "l2_policy": ("working set larger than L2: every pair streams two float64 matrices of windowed distances of up to "
                          "4077 x 4077 x 8 B = 133 MB through HBM (the song arena itself is 38 MB)") if stress else
                         ("the 7 MB song arena fits in L2: this config is the CPU arm's full-set baseline and a parity case, "
                          "not a GPU bench line") if args.config == "covers80" else
                         "inputs larger than L2 (0.45 GB float64 song arena at 10 000 songs, a new tile set every step)"}


# ---------------------------------------------------------------------------- B200 arm
def run_b200(args, rank, world, local):
    import torch
    import torch.distributed as dist
    from acoss_b200 import parallel
    from acoss_b200.serra09 import Serra09

    distributed = world > 1

    def barrier():
        torch.cuda.synchronize()
        if distributed:
            dist.barrier()
        torch.cuda.synchronize()

    # Every rank builds the same seeded song set (in a real run: the feature files on shared storage) and keeps the
    # features of ITS contiguous share for the all-gather; nothing is shipped from rank 0.
    songs, labels = make_workload(args)
    lengths = np.array([s["chroma"].shape[1] for s in songs])
    N = len(songs)
    cpu_base = None
    if rank == 0 and world == 1 and not args.no_cpu_baseline:
        # CPU arm first: its worker processes are forked before this process touches CUDA
        cores = os.cpu_count() or 1
        n = args.cpu_pairs or (2 * cores if args.config == "stress" else max(64, 800 * cores))
        secs = cpu_reference_run(songs, step_pairs(N, n, 10_000, 0, 1), cores)
        cpu_base = {"value": n / secs, "unit": "pairs/s", "cores": cores, "kind": "port",
                    "sample": cpu_sample_desc(n, cores), "seconds": secs}
    if not torch.cuda.is_available():
        raise SystemExit("bench.py --impl b200 needs a CUDA device (there is no CPU fallback)")
    torch.cuda.set_device(local)
    if distributed:
        parallel.init_distributed()
    with contextlib.redirect_stdout(sys.stderr):   # the plugin prints like the reference does; stdout carries ONE JSON line
        alg = Serra09(songs=songs, labels=labels, shortname="bench", do_memmaps=False, do_mfcc=False, device=local, m=M_WIN)
    s0, s1 = parallel.song_shares(N, world)[rank]
    alg.share_songs_allgather(range(s0, s1))   # songs resident on every GPU: shard H2D + all-gather + device-to-device
    ctx = alg._ctx["chroma"][0]
    if args.sweep_gen:
        ctx.set_option("sweep_gen", args.sweep_gen)
    for kv in args.opt:
        name, val = kv.split("=", 1)
        ctx.set_option(name, int(val, 0))
    ts = TileSets(N, lengths, args.pairs_per_step)
    if args.full_job:
        return run_full_job(args, alg, ts, rank, world, barrier)

    # ---- value: C-ABI call, songs resident in HBM, this rank's tiles of the step's tile set
    for s in range(args.warmup):
        ctx.serra09_pairs(ts.rank_pairs(s, world, rank)[0], m=M_WIN, kappa=KAPPA)
    batches = [ts.rank_pairs(args.warmup + s, world, rank) for s in range(args.steps)]
    if args.pair_order != "tiles":
        rng = np.random.default_rng(99)
        batches = [(p[rng.permutation(len(p))] if args.pair_order == "shuffled" else p[::-1], n, im) for p, n, im in batches]
    total_pairs = sum(b[1] for b in batches)
    imbalance = float(np.mean([b[2] for b in batches]))
    sampler = ClockSampler(local)
    ctx.reset_counters()
    stage = {"oti": 0.0, "thresholds": 0.0, "crp": 0.0, "dp": 0.0, "total": 0.0}
    barrier()
    if rank == 0:
        sampler.start()
    t0 = time.perf_counter()
    host = {}
    for pairs, _, _ in batches:
        ctx.serra09_pairs(pairs, m=M_WIN, kappa=KAPPA)   # synchronises its stream before returning
        for k, v in ctx.last_stage_ms().items():
            stage[k] += v
        for k, v in ctx.debug_host_ms().items():
            host[k] = host.get(k, 0.0) + v
    busy = time.perf_counter() - t0                      # this rank's own time, before waiting for the others
    barrier()
    dt = time.perf_counter() - t0
    clocks = sampler.stop() if rank == 0 else None
    launches = ctx.kernel_launches()

    # ---- e2e: the plugin-level distributed pass with HOST buffers, every step: this rank's song shard from pinned
    # host memory -> device, all-gather (device to device), score my tiles, float32 tiles -> rank 0 -> host matrices
    if rank == 0:
        alg.Ds = {k: np.zeros((N, N), np.float32) for k in ("chroma_qmax", "chroma_dmax")}
    e2e_sets = [ts.tile_ids(600 + s) for s in range(args.steps + 2)]
    for ids in e2e_sets[:2]:
        alg.share_songs_allgather(range(s0, s1))
        parallel.score_tiles(alg, ids, TILE, ts.tiles, lengths)
    share_s = gather_s = 0.0
    parts = {"deal_s": 0.0, "score_s": 0.0, "gather_s": 0.0, "gather_wait_s": 0.0, "gather_d2h_s": 0.0, "place_s": 0.0}
    e2e_pairs = 0
    barrier()
    t1 = time.perf_counter()
    for ids in e2e_sets[2:]:
        ta = time.perf_counter()
        alg.share_songs_allgather(range(s0, s1))
        share_s += time.perf_counter() - ta
        info = parallel.score_tiles(alg, ids, TILE, ts.tiles, lengths)
        gather_s += info["gather_place_s"]
        for k in parts:
            parts[k] += info[k]
        e2e_pairs += info["pairs_total"]
    barrier()
    dt_e2e = time.perf_counter() - t1
    return finish_b200(locals())


def add_mfcc(songs, seed=13):
    """Synthetic 13-bin mfcc-like sequences of the songs' lengths (smoothed random walk + noise)."""
    rng = np.random.default_rng(seed)
    out = []
    for sgs in songs:
        L = sgs["chroma"].shape[1]
        fr = rng.normal(size=(L, 13)).cumsum(axis=0).astype(np.float32) * 0.1 + rng.normal(size=(L, 13)).astype(np.float32)
        out.append({"chroma": sgs["chroma"], "gchroma": sgs["gchroma"], "mfcc": np.ascontiguousarray(fr.T)})
    return out


def _cpu_snf_chunk(chunk):
    from oracle import snf_oracle
    out = []
    for i, j in chunk:
        Si, Sj = _CPU_SONGS[int(i)], _CPU_SONGS[int(j)]
        r = snf_oracle.early_snf_pair(Si["chroma"].astype(np.float64), Sj["chroma"].astype(np.float64), Si["gchroma"], Sj["gchroma"],
                                      Si["mfcc"].astype(np.float64), Sj["mfcc"].astype(np.float64), M_WIN, KAPPA, 3)
        out.append(r["qmax_raw"])
    return out


def run_earlysnf(args, rank, world, local):
    """--config earlysnf (BASELINE configs[3]): a step = acoss_snf_pairs over a seeded batch of pairs (fusion of the chroma
    and mfcc kernels + the binarise / Qmax / Dmax tail), the step's pairs split evenly over the ranks; songs resident."""
    import multiprocessing as mp
    import torch
    import torch.distributed as dist
    from acoss_b200 import _native, parallel
    global _CPU_SONGS
    songs, labels = make_workload(args)
    songs = add_mfcc(songs)
    N = len(songs)
    lengths = np.array([s["chroma"].shape[1] for s in songs])
    per_step = args.pairs_per_step if args.pairs_per_step != 1200000 else 384
    cpu_base = None
    if rank == 0 and world == 1 and not args.no_cpu_baseline:
        cores = os.cpu_count() or 1
        n = args.cpu_pairs or 12 * cores   # ~0.8 s per pair and core
        _CPU_SONGS = songs
        pairs = step_pairs(N, n, 10_000, 0, 1)
        with mp.get_context("fork").Pool(cores) as pool:
            pool.map(_cpu_snf_chunk, [pairs[:1]] * cores)
            t0 = time.perf_counter()
            pool.map(_cpu_snf_chunk, [pairs[k:k + 1] for k in range(n)], chunksize=1)
            secs = time.perf_counter() - t0
        cpu_base = {"value": n / secs, "unit": "pairs/s", "cores": cores, "kind": "port", "seconds": secs,
                    "sample": "%d uniformly drawn pairs; oracle/snf_oracle.py (numpy restatement of SimilarityFusion.py, pinned by the "
                              "reference's own output) + the reference's SequenceAlignment.c; %d processes" % (n, cores)}
    if not torch.cuda.is_available():
        raise SystemExit("bench.py needs a CUDA device (there is no CPU fallback)")
    torch.cuda.set_device(local)
    if world > 1:
        parallel.init_distributed()
    ctx = _native.Context(local)
    ctx.upload_songs(songs)
    ctx.upload_aux(np.concatenate([s["mfcc"].T for s in songs], axis=0))

    def my_pairs(step):
        allp = step_pairs(N, per_step, step, 0, 1)
        return allp[rank::world]

    def barrier():
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    for s in range(args.warmup):
        ctx.snf_pairs(my_pairs(s), m=M_WIN, kappa=KAPPA, niters=3)
    batches = [my_pairs(args.warmup + s) for s in range(args.steps)]
    sampler = ClockSampler(local)
    ctx.reset_counters()
    barrier()
    if rank == 0:
        sampler.start()
    t0 = time.perf_counter()
    for b in batches:
        ctx.snf_pairs(b, m=M_WIN, kappa=KAPPA, niters=3)
    barrier()
    dt = time.perf_counter() - t0
    clocks = sampler.stop() if rank == 0 else None
    # algorithmic flops of the cross diffusion: per pair, kind and iteration two products of K n^2 multiply-adds
    flops = 0.0
    for b in batches:
        n = (lengths[b[:, 0]] - M_WIN + 1 + lengths[b[:, 1]] - M_WIN + 1).astype(np.float64)
        flops += float((2 * 2 * 3 * 2.0 * np.floor(KAPPA * n) * n * n).sum())
    tot = torch.tensor([dt, flops, float(ctx.kernel_launches())], dtype=torch.float64, device="cuda")
    if world > 1:
        tmax = tot[:1].clone()
        dist.all_reduce(tmax, op=dist.ReduceOp.MAX)
        dist.all_reduce(tot, op=dist.ReduceOp.SUM)
        tot[0] = tmax[0]
    dt, flops, launches = [float(x) for x in tot]
    if rank == 0:
        sm_max = (clocks or {}).get("sm_max_mhz") or 1965.0
        peak = N_SM * FP64_LANES_PER_SM * 2 * sm_max * 1e6 / 1e12
        total_pairs = per_step * args.steps
        line = {"metric": "song_pairs_per_sec", "value": total_pairs / dt, "unit": "pairs/s", "n_gpus": world, "steps": args.steps,
                "warmup": args.warmup, "ms_per_step": 1e3 * dt / args.steps, "higher_is_better": True, "scaling": "strong",
                "vs_baseline": None, "dtype": "f64", "data": "synthetic",
                "config": {"workload": "%s: %d songs with synthetic 13-bin mfcc, %d pairs per step, m = %d, kappa = %.3f, 3 fusion iterations"
                                       % (CONFIGS[args.config][3], N, per_step, M_WIN, KAPPA), "songs": N, "pairs_per_step_total": per_step,
                           "l2_policy": "inputs larger than L2 (about 32 MB of (M+N)^2 float64 matrices per pair, hundreds of pairs per call)"},
                "clocks": clocks, "gpu_launches": int(launches),
                "e2e": {"value": total_pairs / dt, "unit": "pairs/s", "h2d_bytes_per_step": int(per_step * 8), "d2h_bytes_per_step": int(per_step * 24),
                        "note": "acoss_snf_pairs takes the pair list from the host and returns the scores to the host inside the timed call; the "
                                "songs are resident (uploaded once)"},
                "roofline": {"bound": "tensor", "achieved": flops / dt / 1e12, "peak": peak, "unit": "TFLOP/s", "frac": flops / dt / 1e12 / peak,
                             "traffic": None, "kernel": "k_snf_xst + k_snf_sy (cross diffusion S (P S^T), float64 SIMT gather / axpy)",
                             "note": "algorithmic flops = 2 kinds x 3 iterations x 2 products x 2 K n^2 per pair over the WHOLE step time (all "
                                     "kernels of the call); FP64 peak as for the main path"}}
        if cpu_base is not None:
            line["cpu_baseline"] = cpu_base
        print(json.dumps(line), flush=True)
    ctx.close()
    if world > 1:
        dist.destroy_process_group()


def default_plugin_rate(args, songs, labels, local, n_pairs=100000):
    """The reference's default call computes chroma AND mfcc scores (Serra09.py:94,177-184): pairs/s of
    Serra09(do_mfcc=True).similarity on synthetic 13-bin mfcc-like features of the same lengths (N = 1 only)."""
    from acoss_b200.serra09 import Serra09
    rng = np.random.default_rng(13)
    both = []
    for sgs in songs:
        L = sgs["chroma"].shape[1]
        fr = rng.normal(size=(L, 13)).cumsum(axis=0).astype(np.float32) * 0.1 + rng.normal(size=(L, 13)).astype(np.float32)
        both.append({"chroma": sgs["chroma"], "gchroma": sgs["gchroma"], "mfcc": np.ascontiguousarray(fr.T)})
    with contextlib.redirect_stdout(sys.stderr):
        alg = Serra09(songs=both, labels=labels, shortname="bench_default", do_memmaps=False, do_mfcc=True, device=local)
    alg.preload()
    n = len(songs)
    alg.similarity(step_pairs(n, n_pairs, 900, 0, 1))
    t0 = time.perf_counter()
    sims = alg.similarity(step_pairs(n, n_pairs, 901, 0, 1))
    dt = time.perf_counter() - t0
    out = {"value": n_pairs / dt, "unit": "pairs/s", "pairs": n_pairs, "similarity_types": sorted(sims.keys()),
           "device_ms": {k: round(v["total"], 2) for k, v in alg.last_stage_ms.items()},
           "note": "Serra09(do_mfcc=True).similarity(): chroma + mfcc halves, songs resident, one call; the 13-bin half "
                   "runs on the same k_sweep7 kernels (K padded to 16 for the FP64 tensor-core fragments)"}
    alg.close()
    return out


def run_full_job(args, alg, ts, rank, world, barrier):
    """--full-job: the whole all-pairs matrix once (BASELINE's Target with --config benchmark), through
    parallel.score_tiles in slabs of tiles; prints one JSON line with the wall time (not a bench line)."""
    from acoss_b200 import parallel
    N = alg.N
    if rank == 0:
        alg.Ds = {k: np.zeros((N, N), np.float32) for k in ("chroma_qmax", "chroma_dmax")}
    T = 128
    tiles = parallel.tile_list(N, T)
    lengths = ts.lengths
    slab = max(1, 4_000_000 // (T * T))          # ~4 M pairs per distributed pass
    barrier()
    t0 = time.perf_counter()
    pairs = 0
    score_s = gather_s = 0.0
    imb = []
    for a in range(0, len(tiles), slab):
        info = parallel.score_tiles(alg, range(a, min(a + slab, len(tiles))), T, tiles, lengths)
        pairs += info["pairs_total"]
        score_s += info["score_s"]
        gather_s += info["gather_place_s"]
        for k in parts:
            parts[k] += info[k]
        imb.append(info["lpt_imbalance"])
    barrier()
    wall = time.perf_counter() - t0
    if rank == 0:
        for k in alg.Ds:
            alg.Ds[k] += alg.Ds[k].T
        q = alg.Ds["chroma_qmax"]
        print(json.dumps({"full_job": True, "config": workload_config(args), "n_gpus": world, "pairs": pairs, "wall_s": wall,
                          "pairs_per_sec": pairs / wall, "rank0_score_s": score_s, "rank0_gather_place_s": gather_s,
                          "lpt_imbalance_mean": float(np.mean(imb)), "lpt_imbalance_max": float(np.max(imb)),
                          "symmetric": bool(np.array_equal(q, q.T)), "nonzero_fraction": float((q > 0).mean()),
                          "checksum_qmax": float(q.sum(dtype=np.float64))}), flush=True)
    alg.close()


def finish_b200(v):
    import torch
    import torch.distributed as dist
    args, rank, world, local = v["args"], v["rank"], v["world"], v["local"]
    alg, ctx, lengths, stage, batches = v["alg"], v["ctx"], v["lengths"], v["stage"], v["batches"]
    tmax = torch.tensor([v["dt"], v["dt_e2e"], v["share_s"], v["gather_s"], v["busy"]], dtype=torch.float64, device="cuda")
    tmin = torch.tensor([v["busy"]], dtype=torch.float64, device="cuda")
    my = [float(sum(x)) for x in zip(*[pair_work(lengths, b[0]) for b in batches])]
    work = torch.tensor(my + [float(v["launches"])], dtype=torch.float64, device="cuda")
    if v["distributed"]:
        dist.all_reduce(tmax, op=dist.ReduceOp.MAX)
        dist.all_reduce(tmin, op=dist.ReduceOp.MIN)
        dist.all_reduce(work, op=dist.ReduceOp.SUM)
    dt, dt_e2e, share_s, gather_s, busy_max = [float(x) for x in tmax]
    cells, qcells, abytes, launches = [float(x) for x in work]
    if rank != 0:
        alg.close()
        return
    steps = args.steps
    value = v["total_pairs"] / dt
    e2e_value = v["e2e_pairs"] / dt_e2e
    pairs_step = v["total_pairs"] / steps
    h2d = int(sum(a.nbytes for a in alg._shards["chroma"][1] if a is not None) * world + pairs_step * 104)
    d2h = int(pairs_step * 8 * (2 if world > 1 else 1))
    peaks = {}
    try:
        peaks = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
    except Exception:
        pass
    my_cells, my_q, my_bytes = my
    sweep_s = (stage["thresholds"] + stage["crp"]) / 1e3
    clocks = v["clocks"]
    sm_max = (clocks or {}).get("sm_max_mhz") or peaks.get("sm_max_mhz") or 1965.0
    fp64_peak = N_SM * FP64_LANES_PER_SM * 2 * sm_max * 1e6 / 1e12
    achieved_tf = 26.0 * my_cells / sweep_s / 1e12 if sweep_s > 0 else None
    hbm_peak = peaks.get("hbm_gbs", 6650.0)
    traffic = None
    try:  # DRAM bytes per pair of the sweep launches, from the ncu --set full capture of the same command
        if args.config == "stress":
            raise KeyError("no capture for this config")
        tj = json.load(open(os.path.join(ROOT, "profiles", "r02_sweep_traffic.json")))
        traffic = tj["dram_bytes_per_pair"] * (len(batches[0][0]) if batches else 0)
    except Exception:
        pass
    dp_s = stage["dp"] / 1e3
    # integer min / max / add instructions (VIADDMNMX, VIMNMX3, their .S16x2 forms) issue at 64 lanes per clock per SM
    # on the B200 (profiles/exp/ubench/dpx_rate_b200.txt), half of what SURVEY 8d assumed; the packed kernel spends
    # 5 of them on a cell's Qmax AND Dmax update, i.e. 2.5 lane-ops per cell update
    int_peak_gcups = N_SM * 64 * sm_max * 1e6 / 2.5 / 1e9
    dp_gcups = (2.0 * my_q / dp_s / 1e9) if dp_s > 0 else None
    line = {
        "metric": "song_pairs_per_sec", "value": value, "unit": "pairs/s", "n_gpus": world, "steps": steps,
        "warmup": args.warmup, "ms_per_step": 1e3 * dt / steps, "higher_is_better": True, "scaling": "strong",
        "vs_baseline": None, "dtype": "f64", "data": "synthetic", "config": workload_config(args),
        "pairs_per_step": pairs_step,
        "qmax_gcups_whole_path": qcells / dt / 1e9,
        "qmax_dmax_gcups_dp_kernel": dp_gcups,
        "cells_per_sec": cells / dt,
        "device_ms_per_step": {k: x / steps for k, x in stage.items()},
        "host_ms_per_step": {k: x / steps for k, x in v["host"].items()},
        "clocks": clocks,
        "lpt_imbalance": v["imbalance"],
        "rank_busy_ms_per_step": {"max": 1e3 * busy_max / steps, "min": 1e3 * float(tmin[0]) / steps},
        "e2e": {"value": e2e_value, "unit": "pairs/s", "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": d2h,
                "ms_per_step": 1e3 * dt_e2e / steps, "share_songs_ms_per_step": 1e3 * share_s / steps,
                "gather_place_ms_per_step": 1e3 * gather_s / steps,
                "rank0_ms_per_step": {k[:-2]: 1e3 * x / steps for k, x in v["parts"].items()}},
        "gpu_launches": int(launches),
        "roofline": {"bound": "tensor", "achieved": achieved_tf, "peak": fp64_peak, "unit": "TFLOP/s",
                     "frac": (achieved_tf / fp64_peak) if achieved_tf else None, "traffic": traffic,
                     "kernel": ("k_sweep_tile + k_select_rows_cta + k_emit_crp (tiled long-song sweep, float64 SIMT, windowed "
                                "distances through HBM)") if args.config == "stress" else
                               "k_sweep7 (FP64 tensor-core dot products + integer window sums + chained row selection: "
                               "separators of one orientation, packed CRP of the other); k_sweep4 / k_sweep_fast / the tiled kernels only "
                               "for column songs over 1015 frames",
                     "note": "FP64 roofline, NOT the bf16 figure of MEASURED_PEAKS.json (which has no FP64 entry): "
                             "peak = 148 SM x 64 FP64 FMA/cycle x 2 x sm_max_mhz, the rate measured for both DMMA "
                             "(mma.sync.m8n8k4.f64, profiles/exp/ubench/dmma_rate.cu) and DFMA; algorithmic 26 flop/cell "
                             "counted ONCE although every cell is computed twice (once per orientation); achieved = flops "
                             "of rank 0's sweep launches of the timed steps / their summed device time (CUDA events on "
                             "the launching stream); traffic = DRAM bytes of the sweep launches for one step's pairs of "
                             "rank 0 (ncu dram__bytes_read.sum + dram__bytes_write.sum, profiles/r02_sweep_traffic.json)"},
        "roofline_dp": {"bound": "int-alu", "achieved": dp_gcups, "peak": int_peak_gcups, "unit": "GCUPS",
                        "frac": (dp_gcups / int_peak_gcups) if dp_gcups else None, "kernel": "k_dp16<4> (Qmax + Dmax cell updates on packed s16x2 lanes)",
                        "note": "peak = 148 SM x 64 integer min/max lanes per clock (measured: profiles/exp/ubench/dpx_rate_b200.txt; "
                                "SURVEY.md 8d assumed 128) x sm_max_mhz / 2.5 lane-ops per cell update (5 VIADDMNMX.S16x2 per cell for "
                                "both tables)"},
        "roofline_hbm": {"bound": "hbm", "achieved": (my_bytes / sweep_s / 1e9) if sweep_s > 0 else None, "peak": hbm_peak,
                         "unit": "GB/s", "frac": (my_bytes / sweep_s / 1e9 / hbm_peak) if sweep_s > 0 else None,
                         "note": "compulsory bytes only (both songs' frames + scores); the path is compute-bound by design, "
                                 "a high number here would mean the CRP round-trips (peak: measured copy bandwidth)"},
    }
    if v["cpu_base"] is not None:
        line["cpu_baseline"] = v["cpu_base"]
    if world == 1 and args.config != "stress":
        try:
            line["e2e_default_plugin"] = default_plugin_rate(args, v["songs"], v["labels"], local)
        except Exception as e:  # the headline must not depend on this extra
            line["e2e_default_plugin"] = {"error": repr(e)}
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
    if args.config == "earlysnf":
        run_earlysnf(args, rank, world, local)
        return
    run_b200(args, rank, world, local)
    if world > 1:
        import torch.distributed as dist
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
