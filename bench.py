#!/usr/bin/env python
"""bench.py -- headline benchmark of the population self-replication hot path.

  python bench.py [--gpus N] [--steps K] [--warmup W] [--impl reference]

Workload (config.workload = "ww22_self_application"): per GPU, 2^24 WeightwiseNeuralNetwork(width=2, depth=2)
nets with Keras-default (glorot-uniform) weights, 100 unconditional self-applications each -- the shape of
BASELINE.json's configs[0]/[4] (code/setups/applying-fixpoints.py: 100 x self_attack per net) at population
scale.  One "step" = one pass of the kernel over one arena of 2^24 nets = 1.68e9
self-applications.  Every timed step works on its own fresh arena (0.94 GB each, far larger than the 126 MB
L2), all resident in HBM before the timed region starts.

  value    whole-job self-applications/s, device-timed (CUDA events), max over ranks
  e2e      same metric through the host API with HOST inputs: pinned-host -> device copy of the step's
           14-float records, kernel + fused classification, final weights and class counters back to the host,
           three arenas in flight
  roofline the kernel against the FP32 issue roofline (bound "fp32"): achieved = 392 algorithmic FLOP per
           self-application / kernel time; peak = an FFMA-only probe kernel measured live on the same GPU
  cpu_baseline  the oracle's SIMD/OpenMP WW(2,2) path on the box's host cores, on a bounded sample

--impl reference: times the CPU implementation alone (the oracle port -- the reference's own Keras path needs
TensorFlow 1.x, which is not installable here or on the box), rank 0 only.
"""
import argparse
import json
import os
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
PKG = os.path.join(ROOT, "self-replicating-neural-networks_b200")
for _p in (ROOT, PKG):
    if _p not in sys.path:
        sys.path.insert(0, _p)

import numpy as np  # noqa: E402

N_NETS = 1 << 24
N_STEPS = 100
FLOP_PER_SELF_APP = 392.0       # 14 points x 14 MAC x 2
FP32_INSTR_PER_SELF_APP = 210.0  # FMA-pipe lane-ops as issued: 7 point pairs x 15 packed instructions x 2 lanes
BYTES_PER_NET = 112.0           # 56 B read + 56 B write per resident episode
# dram__bytes_read.sum + dram__bytes_write.sum of one ncu --set full capture of this kernel at 2^24 nets
# (profiles/r02_k1_self_apply_fixed_ncu_summary.csv, unchanged since round 1): 0.941 GB + 0.880 GB per launch
NCU_TRAFFIC_BYTES_PER_NET = (940.291584e6 + 880.501760e6) / (1 << 24)
# the same for one generation of the 10 M-particle soup = soup_lists_kernel + soup_sync_kernel + soup_deferred_learners_kernel
# (profiles/r02_soup_sync_10m_ncu_summary.csv: 0.059 + 1.407 + 0.411 GB read + written)
SOUP_10M_NCU_TRAFFIC = 1.877e9
SOUP_10M_NCU_SOURCE = ("ncu dram bytes of the three launches of one generation at 10 M particles (lists 0.06 GB, tile kernel "
                       "0.86 + 0.55 GB, deferred learners 0.34 + 0.07 GB), profiles/r02_soup_sync_10m_ncu_summary.csv")
METRIC = "net self-applications/sec"
UNIT = "self-applications/s"


class ClockSampler:
    """SM clock and throttle reasons DURING the timed region (B200_PROFILING.md): NVML polled every 5 ms from a
    thread (falls back to `nvidia-smi -lms` when the NVML binding is unavailable)."""
    REASONS = (("hw_slowdown", 0x8), ("sw_thermal_slowdown", 0x20), ("hw_thermal_slowdown", 0x40),
               ("sw_power_cap", 0x4))

    def __init__(self, index):
        self.index, self.sm, self.mx, self.reasons, self.power = index, [], None, set(), []
        self._stop, self._thread, self.source = threading.Event(), None, None

    def _poll_nvml(self, nv, h):
        while not self._stop.is_set():
            try:
                self.sm.append(float(nv.nvmlDeviceGetClockInfo(h, nv.NVML_CLOCK_SM)))
                mask = int(nv.nvmlDeviceGetCurrentClocksThrottleReasons(h))
                for name, bit in self.REASONS:
                    if mask & bit:
                        self.reasons.add(name)
                self.power.append(nv.nvmlDeviceGetPowerUsage(h) / 1000.0)
            except Exception:
                pass
            time.sleep(0.005)

    def _poll_smi(self):
        q = ("clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
             "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")
        proc = subprocess.Popen(["nvidia-smi", "-i", str(self.index), "--query-gpu=" + q, "--format=csv,noheader,nounits",
                                 "-lms", "50"], stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
        names = ("hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap")
        while not self._stop.is_set():
            line = proc.stdout.readline()
            if not line:
                break
            c = [x.strip() for x in line.split(",")]
            try:
                self.sm.append(float(c[0])); self.mx = float(c[1])
            except (ValueError, IndexError):
                continue
            for name, val in zip(names, c[2:]):
                if val.lower().startswith("active"):
                    self.reasons.add(name)
        proc.terminate()

    def start(self):
        try:
            import pynvml as nv
            nv.nvmlInit()
            uuid = None
            try:
                import torch
                uuid = str(torch.cuda.get_device_properties(self.index).uuid)
            except Exception:
                pass
            h = None
            if uuid:
                for cand in (uuid, "GPU-" + uuid):
                    try:
                        h = nv.nvmlDeviceGetHandleByUUID(cand.encode() if isinstance(cand, str) else cand)
                        break
                    except Exception:
                        h = None
            if h is None:
                h = nv.nvmlDeviceGetHandleByIndex(self.index)
            self.mx = float(nv.nvmlDeviceGetMaxClockInfo(h, nv.NVML_CLOCK_SM))
            self.source = "nvml"
            self._thread = threading.Thread(target=self._poll_nvml, args=(nv, h), daemon=True)
        except Exception:
            self.source = "nvidia-smi"
            self._thread = threading.Thread(target=self._poll_smi, daemon=True)
        self._thread.start()

    def stop(self):
        self._stop.set()
        if self._thread:
            self._thread.join(timeout=2.0)
        return {"sm_mhz": float(np.median(self.sm)) if self.sm else None, "sm_max_mhz": self.mx,
                "reasons": sorted(self.reasons), "samples": len(self.sm), "source": self.source,
                "power_w_max": max(self.power) if self.power else None}


def cpu_port_rate(n_nets, seconds):
    """Oracle WW(2,2) SIMD/OpenMP path, repeated for ~`seconds`; returns (self-applications/s, cores, sample)."""
    import oracle
    rng = np.random.default_rng(0)
    lim = np.array([1.0] * 8 + [np.sqrt(1.5)] * 4 + [np.sqrt(2.0)] * 2)
    w0 = (rng.uniform(-1, 1, (n_nets, 14)) * lim).astype(np.float32)
    oracle.ww22_self_apply_fast(w0.copy(), 2)  # warm-up (thread pool, page faults)
    done, t0 = 0, time.perf_counter()
    while True:
        w = w0.copy()
        t1 = time.perf_counter()
        oracle.ww22_self_apply_fast(w, N_STEPS)
        done += 1
        el = time.perf_counter() - t0
        if el >= seconds:
            break
    dt = time.perf_counter() - t0
    return n_nets * N_STEPS * done / dt, oracle.num_threads(), "%d x (2^%d nets x %d steps)" % (
        done, int(np.log2(n_nets)), N_STEPS)


def headline_config(n):
    """The workload both arms are timed on (identical dict in the B200 line and the --impl reference line)."""
    return {"workload": "ww22_self_application", "nets_per_gpu": n, "self_applications_per_net": N_STEPS,
            "net": "WeightwiseNeuralNetwork(width=2, depth=2)", "init": "glorot_uniform",
            "l2": "inputs larger than L2 (0.94 GB arena per step, fresh arena each step)",
            "parallelism": "population sharded by net index, no data-path collective"}


def bind_to_numa_node(rank):
    """Pin this process to the CPUs of NUMA node (rank mod #nodes); returns the CPU list or None."""
    try:
        nodes = sorted(int(d[4:]) for d in os.listdir("/sys/devices/system/node") if d.startswith("node") and d[4:].isdigit())
        node = nodes[rank % len(nodes)]
        cpus = []
        for part in open("/sys/devices/system/node/node%d/cpulist" % node).read().strip().split(","):
            lo, _, hi = part.partition("-")
            cpus.extend(range(int(lo), int(hi or lo) + 1))
        allowed = sorted(set(cpus) & set(os.sched_getaffinity(0)))
        if not allowed:
            return None
        os.sched_setaffinity(0, allowed)
        return allowed
    except Exception:
        return None


def host_cores():
    try:
        return len(os.sched_getaffinity(0))
    except AttributeError:
        return os.cpu_count() or 1


def run_reference(args, rank):
    """The reference arm: the path's CPU implementation on every host core of the box.  The reference's own Keras /
    TensorFlow-1.x code cannot be installed (no oracle/_ref), so this is the oracle's SIMD + OpenMP port of
    WeightwiseNeuralNetwork.apply_to_weights (same pinned arithmetic).  One step = the FULL 2^24-net arena x 100
    self-applications, exactly the B200 arm's step.  Under torchrun only rank 0 works -- with all cores: the
    launcher's OMP_NUM_THREADS=1 is overridden."""
    if rank != 0:
        return
    import oracle
    cores = host_cores()
    oracle.set_threads(cores)
    n = args.nets
    rng = np.random.default_rng(0)
    lim = np.array([1.0] * 8 + [np.sqrt(1.5)] * 4 + [np.sqrt(2.0)] * 2)
    w0 = np.empty((n, 14), np.float32)
    for lo in range(0, n, 1 << 20):
        hi = min(n, lo + (1 << 20))
        w0[lo:hi] = (rng.uniform(-1, 1, (hi - lo, 14)) * lim).astype(np.float32)
    w = np.empty_like(w0)
    times = []
    for s in range(args.warmup + args.steps):
        np.copyto(w, w0)                       # fresh inputs every step, outside the timed region
        t0 = time.perf_counter()
        oracle.ww22_self_apply_fast(w, N_STEPS)
        dt = time.perf_counter() - t0
        if s >= args.warmup:
            times.append(dt)
    total = float(np.sum(times))
    val = n * N_STEPS * args.steps / total
    sample = "every step = all 2^%d nets x %d self-applications (nothing sampled down)" % (int(np.log2(n)), N_STEPS)
    print(json.dumps({
        "impl": "reference", "metric": METRIC, "value": val, "unit": UNIT, "n_gpus": args.gpus, "steps": args.steps,
        "warmup": args.warmup, "ms_per_step": total / args.steps * 1e3, "higher_is_better": True, "scaling": "weak",
        "vs_baseline": None, "dtype": "f32", "data": "synthetic", "config": headline_config(n),
        "cpu_baseline": {"value": val, "unit": UNIT, "cores": oracle.num_threads(), "kind": "port", "sample": sample},
        "e2e": {"value": val, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "note": "reference Keras/TensorFlow-1.x path unavailable on this box; timed the oracle port (same arithmetic, "
                "AVX-512/AVX2 clones + OpenMP over nets) on all host cores; at N > 1 the value is still ONE host's "
                "CPU rate -- it does not grow with the number of GPUs"}), flush=True)


def _timed(torch, fn, reps=1):
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    torch.cuda.synchronize()
    e0.record()
    for _ in range(reps):
        fn()
    e1.record()
    e1.synchronize()
    return e0.elapsed_time(e1) * 1e-3 / reps


def _hbm_peak():
    try:
        return float(json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))["hbm_gbs"]), "MEASURED_PEAKS.json"
    except Exception:
        return 6650.0, "fallback (B200_PROFILING.md)"


def secondary_legs(eng, torch, dist, rank, world, dev, probe):
    """The other half of BASELINE.json's metric (soup generations/s) and configs 2-5, device-timed, max over ranks.
    Reported as extra keys of the JSON line; the headline `value` stays net self-applications/s.
    Every N: self-training (config 2, weak-scaled), the mixed 10 M soup (config 4), the KFV sweep (config 5, weak-scaled).
    N = 1 only: the 100k soups of config 3 (a one-GPU config), size sweep, early-stop variant, the class-API facade."""
    from soup import PopulationSoup, evolve_together
    out = {}

    def sync_max(x):
        t = torch.tensor([x], dtype=torch.float64, device=dev)
        if world > 1:
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return t.item()

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def timed(fn):
        barrier()
        return sync_max(_timed(torch, fn))

    # ---- configs[1]: self-training.  Per GPU 2^20 WW(2,2) nets x 100 epochs (Philox sample orders), weak-scaled;
    #      plus the config as written (10k WW + 10k AGG(4) x 1000 epochs) on rank 0's GPU
    dw, da = eng.make_desc(eng.KIND_WW, 2, 2), eng.make_desc(eng.KIND_AGG, 2, 2, 4)
    big = eng.Population(dw, 1 << 20).init(seed=5, slot0=rank << 20)
    big.train(2, perm_mode=eng.PERM_PHILOX, seed=2, slot0=rank << 20, want_loss=False)
    dtb = timed(lambda: big.train(100, perm_mode=eng.PERM_PHILOX, seed=2, slot0=rank << 20, want_loss=False))
    pw, pa = eng.Population(dw, 10_000).init(seed=1), eng.Population(da, 10_000).init(seed=1)
    pw.train(10, perm_mode=eng.PERM_PHILOX, seed=2, want_loss=False)
    dt = _timed(torch, lambda: (pw.train(1000, perm_mode=eng.PERM_PHILOX, seed=2, want_loss=False),
                                pa.train(1000, want_loss=False)))
    tf = (1 << 20) * 100 * 1148.0 / dtb / 1e12           # per GPU
    out["self_train"] = {"metric": "net self-train epochs/sec", "value": world * (1 << 20) * 100 / dtb, "unit": "epochs/s",
                         "config": {"workload": "ww22_self_train", "nets_per_gpu": 1 << 20, "epochs": 100, "perm": "philox",
                                    "scaling": "weak"},
                         "config_10k_ww_10k_agg_1000_epochs_s": dt,
                         "roofline": {"bound": "fp32", "achieved": tf, "peak": probe * 2 / 1e12, "unit": "TFLOP/s",
                                      "frac": tf / (probe * 2 / 1e12), "kernel": "train_kernel<WW(2,2)>",
                                      "algorithmic_flop": "1148 per epoch (14 samples x 82)",
                                      "fma_pipe_lane_ops": "924 per epoch in the pinned order (update = 3 roundings per weight)",
                                      "fma_pipe_frac": (1 << 20) * 100 * 924.0 / dtb / probe},
                         "classes_10k_ww": dict(zip(eng.CLS_NAMES, pw.classify(1e-4)[1].cpu().tolist())),
                         "classes_10k_agg": dict(zip(eng.CLS_NAMES, pa.classify(1e-4)[1].cpu().tolist()))}
    del big, pw, pa

    # ---- configs[4]: known-fixpoint variation sweep, per GPU 100 M perturbed identity fixpoints per noise level
    #      1e0 .. 1e-9, <= 100 self-applications each, weak-scaled (one histogram all-reduce at the end)
    basefix = np.zeros(14, np.float32)
    basefix[[0, 8, 12]] = 1.0
    ref_y = [3.63, 5.02, 6.46, 8.04, 9.61, 11.23, 12.99, 14.58, 21.95, 26.45]
    ref_z = [0, 0, 0, 0, 0.04, 1.38, 3.23, 4.84, 11.91, 16.47]
    nt, scale, t_kfv = 100_000_000, 1.0, 0.0
    eng.kfv(dw, basefix, nt, 1.0, 0, 100, 1e-4, seed=11, slot0=rank * nt, want_times=False)
    hists = []
    for level in range(10):
        res = {}
        t_kfv += timed(lambda: res.update(eng.kfv(dw, basefix, nt, scale, level, 100, 1e-4, seed=11, slot0=rank * nt,
                                                  want_times=False)))
        hists.append(res["hist"])
        scale /= 10.0
    h = torch.stack(hists).to(torch.float64)
    if world > 1:
        dist.all_reduce(h)
    h = h.cpu().numpy()
    steps = np.arange(101)
    ys = [float((h[l, 0] * steps).sum() / (world * nt)) for l in range(10)]
    zs = [float((h[l, 1] * steps).sum() / (world * nt)) for l in range(10)]
    steps_total = float(sum((h[l, 0] * steps).sum() for l in range(10))) + 10.0 * world * nt
    out["kfv_sweep"] = {"metric": "known-fixpoint variation trials/sec", "value": 10 * world * nt / t_kfv, "unit": "trials/s",
                        "config": {"workload": "kfv_sweep", "trials_per_level_per_gpu": nt, "levels": "1e0..1e-9",
                                   "max_steps": 100, "epsilon": 1e-4, "scaling": "weak"},
                        "self_applications_per_s": steps_total / t_kfv,
                        "mean_steps_to_vergence": ys, "reference_log_steps_to_vergence": ref_y,
                        "mean_steps_as_fixpoint": zs, "reference_log_steps_as_fixpoint": ref_z}

    # ---- configs[3]: mixed soup of 10 M nets (one third each WW / AGG / RNN, partners within kind), learn_from + attack.
    #      N = 1: one owner.  N > 1: sharded by particle index; partner records read from the owning GPU inside the
    #      generation kernel ("peer") and, for comparison, the per-generation NCCL all-gather the north_star names.
    third = (10_000_000 // 3) // 8 * 8 // world * world
    kw = dict(attacking_rate=0.1, learn_from_rate=0.1, train=0, learn_from_severity=1, epsilon=1e-4)
    kinds = ((eng.KIND_WW, 0), (eng.KIND_AGG, 4), (eng.KIND_RNN, 0))
    gens, res, host = 20, {}, {}
    for exchange in (("allgather", "peer_unchained", "peer") if world > 1 else ("auto",)):
        soups = [PopulationSoup(k, third, aggregates=a, seed=5 + k, exchange=exchange, **kw).seed_particles()
                 for k, a in kinds]
        evolve_together(soups, 2)
        res[soups[0].exchange] = timed(lambda: evolve_together(soups, gens)) / gens
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        evolve_together(soups, gens)                 # host time to ENQUEUE a generation (no device sync)
        host[soups[0].exchange] = sync_max((time.perf_counter() - t0) / gens)
        torch.cuda.synchronize()
        del soups
        torch.cuda.empty_cache()
    best = min(res, key=res.get)
    soup = {"metric": "soup generations/sec", "value": 1.0 / res[best], "unit": "generations/s",
            "config": dict(workload="mixed_soup_10m", particles=3 * third, kinds="WW/AGG/RNN thirds", ranks=world,
                           generations=gens, exchange=best, scaling="strong", mode="snapshot-synchronous", **kw),
            "generations_per_s_by_exchange": {k: 1.0 / v for k, v in res.items()},
            "host_enqueue_ms_per_generation": {k: v * 1e3 for k, v in host.items()},
            "allgather_bytes_per_generation_per_gpu": sum(third * nw * 4 for nw in (14, 20, 17)) if world > 1 else 0}
    if world > 1:
        # parity inside the bench: a sharded soup must equal the same soup owned by ONE process, bit for bit
        small = 3 * 4096 * world
        same = True
        for k, a in kinds:
            sh = PopulationSoup(k, small, aggregates=a, seed=77, exchange="auto", **kw).seed_particles().evolve(3)
            one = PopulationSoup(k, small, aggregates=a, seed=77, sharded=False, **kw).seed_particles().evolve(3)
            mine = one.w[sh.slot0:sh.slot0 + sh.n_local].contiguous().view(torch.int32)
            same = same and bool(torch.equal(sh.w.contiguous().view(torch.int32), mine))
            del sh, one
        flag = torch.tensor([1 if same else 0], dtype=torch.int32, device=dev)
        dist.all_reduce(flag, op=dist.ReduceOp.MIN)
        soup["sharded_equals_single_owner_bitwise"] = bool(flag.item())
        assert flag.item() == 1, "sharded soup differs from the single-owner soup"
    out["soup_mixed_10m"] = soup
    if world > 1:
        out["soup"] = soup              # key the round-1 driver records read
        return out

    # ---- N = 1 only below
    # configs[2]: soup of 100k WW, attack 0.1 + self-train 1 epoch, respawn of divergent / zero nets, 1000 generations
    kw = dict(attacking_rate=0.1, learn_from_rate=-1, train=1, remove_divergent=True, remove_zero=True, epsilon=1e-4)
    s = PopulationSoup(eng.KIND_WW, 100_000, seed=3, **kw).seed_particles()
    s.evolve(50)
    dt = _timed(torch, lambda: s.evolve(1000))
    out["soup"] = {"metric": "soup generations/sec", "value": 1000 / dt, "unit": "generations/s",
                   "config": dict(workload="soup_100k_ww", particles=100_000, generations=1000, mode="snapshot-synchronous", **kw),
                   "particle_generations_per_s": 100_000 * 1000 / dt, "classes_after": s.count()}
    # the same soup with the reference's exact sequential semantics (dependency-levelled wavefront, bit-identical
    # to Soup.evolve's particle-by-particle sweep)
    s = PopulationSoup(eng.KIND_WW, 100_000, seed=3, order="exact", **kw).seed_particles()
    s.evolve(10)
    dt = _timed(torch, lambda: s.evolve(200))
    out["soup_exact_order"] = {"metric": "soup generations/sec", "value": 200 / dt, "unit": "generations/s",
                               "config": dict(workload="soup_100k_ww_exact_order", particles=100_000, generations=200,
                                              mode="exact sequential order (wavefront)", **kw),
                               "rounds_last_generation": s.rounds, "classes_after": s.count()}
    # 10 M particles, train=0 (configs[3]'s per-GPU regime): one HBM pass per generation, both orders
    kw0 = dict(attacking_rate=0.1, learn_from_rate=0.1, train=0, remove_divergent=True, remove_zero=True, epsilon=1e-4)
    hbm_peak, peak_src = _hbm_peak()
    s = PopulationSoup(eng.KIND_WW, 10_000_000, seed=4, **kw0).seed_particles()
    s.evolve(4)
    dt = _timed(torch, lambda: s.evolve(20))
    ach = 10_000_000 * 130.0 * 20 / dt / 1e9
    out["soup_10m"] = {"metric": "soup generations/sec", "value": 20 / dt, "unit": "generations/s",
                       "config": dict(workload="soup_10m_ww_train0", particles=10_000_000, generations=20, **kw0),
                       "roofline": {"bound": "hbm", "achieved": ach, "peak": hbm_peak, "unit": "GB/s", "frac": ach / hbm_peak,
                                    "traffic": SOUP_10M_NCU_TRAFFIC, "peak_source": peak_src,
                                    "kernel": "soup_sync_kernel<WW(2,2)> (+ lists and deferred-learners launches of the generation)",
                                    "algorithmic_bytes": "130 B per particle-generation (112 B own record in/out + 0.2 x 56 B "
                                                         "partner records + list heads)",
                                    "traffic_source": SOUP_10M_NCU_SOURCE}}
    del s
    s = PopulationSoup(eng.KIND_WW, 10_000_000, seed=4, order="exact", **kw0).seed_particles()
    s.evolve(2)
    dt = _timed(torch, lambda: s.evolve(10))
    out["soup_10m_exact_order"] = {"metric": "soup generations/sec", "value": 10 / dt, "unit": "generations/s",
                                   "config": dict(workload="soup_10m_ww_train0_exact_order", particles=10_000_000,
                                                  generations=10, mode="exact sequential order (wavefront)", **kw0),
                                   "rounds_last_generation": s.rounds}
    del s
    # configs[2] variant of soup_trajectorys.py:22-24: 30 self-train epochs per particle and generation
    kw30 = dict(attacking_rate=0.1, learn_from_rate=-1, train=30, remove_divergent=True, remove_zero=True, epsilon=1e-4)
    s = PopulationSoup(eng.KIND_WW, 100_000, seed=3, **kw30).seed_particles()
    s.evolve(5)
    dt = _timed(torch, lambda: s.evolve(100))
    out["soup_train30"] = {"metric": "soup generations/sec", "value": 100 / dt, "unit": "generations/s",
                           "config": dict(workload="soup_100k_ww_train30", particles=100_000, generations=100, **kw30),
                           "self_train_epochs_per_s": 100_000 * 30 * 100 / dt, "classes_after": s.count()}
    del s
    # configs[0/1] side figures: the size sweep 2^20 .. 2^27 nets (fresh glorot nets, 100 unconditional steps) and the
    # early-stop variant (run_net's stop rule, experiment.py:70-77: stop at divergence or fixpoint), which executes
    # fewer applications per net; both device-timed, one launch each
    sweep = {}
    for lg in range(20, 28):
        pop = eng.Population(dw, 1 << lg).init(seed=lg)
        if lg == 20:
            pop.self_apply(N_STEPS)
            pop.init(seed=lg)
        sweep["2^%d" % lg] = (1 << lg) * N_STEPS / _timed(torch, lambda: pop.self_apply(N_STEPS))
        del pop
    out["size_sweep"] = {"metric": METRIC, "unit": UNIT, "values": sweep}
    es = {}
    for eps in (1e-14, 1e-4):
        pop = eng.Population(dw, N_NETS).init(seed=0)
        res = {}
        dt = _timed(torch, lambda: res.update(pop.self_apply(N_STEPS, mode=eng.RUN_UNTIL, eps=eps, want_steps=True,
                                                             want_counters=True)))
        executed = int(res["steps"].sum(dtype=torch.int64).item())
        es["eps=%g" % eps] = {"nets_per_s": N_NETS / dt, "executed_self_applications_per_s": executed / dt,
                              "mean_steps": executed / N_NETS,
                              "classes": dict(zip(eng.CLS_NAMES, res["counters"].cpu().tolist()))}
        del pop
    out["early_stop"] = {"workload": "ww22_run_net_until_vergence", "nets": N_NETS, "step_limit": N_STEPS, "by_epsilon": es}
    # configs[0] exactly as the reference writes it: 100 nets x 100 self_attack() + count() through the CLASS API (one
    # Python object and one tiny launch per call), next to the reference-structured NumPy twin on one host core
    try:
        sys.path.insert(0, os.path.join(ROOT, "tools"))
        import facade_bench
        from oracle import numpy_twin
        fac = facade_bench.run()
        fac["numpy_twin_1core_self_applications_per_s"] = numpy_twin.self_applications_per_second()
        fac["note"] = ("per-object path: host-bound by construction (self_attack() calls of a net are deferred into one "
                       "launch, then one launch + one device read per net for count); populations go through Soup / "
                       "srnn_engine")
        out["facade"] = fac
    except Exception as exc:        # the facade leg must never take the headline down
        out["facade"] = {"error": repr(exc)}
    return out


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--nets", type=int, default=N_NETS, help="nets per GPU (default 2^24)")
    ap.add_argument("--cpu-seconds", type=float, default=10.0)
    ap.add_argument("--no-legs", action="store_true", help="headline + e2e only (profiling runs)")
    args = ap.parse_args()
    args.warmup = max(args.warmup, 3) if args.impl == "b200" else args.warmup
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    if args.impl == "reference":
        run_reference(args, rank)
        return

    import torch
    import torch.distributed as dist
    import srnn_engine as eng
    torch.cuda.set_device(local_rank)
    dev = torch.device("cuda", local_rank)
    host_cpus, host_policy = None, "unbound"
    if world > 1:
        # one process per GPU.  Where the pinned staging buffers of the e2e leg live decides its scaling: they are
        # first-touched by this thread, so the thread is placed before they are allocated.  SRNN_BENCH_HOST =
        #   local      the GPU's own NUMA node (NVML affinity)
        #   spread     rank r on NUMA node r mod (number of nodes): when every GPU hangs off the same node, that node's
        #              memory is what 8 ranks x 1.9 GB per step saturate; half the traffic then goes through the other socket
        #   unbound    wherever the launcher put the process
        # (N = 1 keeps every core: the cpu_baseline leg uses them all)
        host_policy = os.environ.get("SRNN_BENCH_HOST", "spread")
        if host_policy == "local":
            host_cpus = eng.bind_host_thread_to_gpu(local_rank)
        elif host_policy == "spread":
            host_cpus = bind_to_numa_node(local_rank)
        dist.init_process_group("nccl", device_id=dev)
    n = args.nets
    K, W = args.steps, args.warmup
    desc = eng.make_desc(eng.KIND_WW, 2, 2)

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    # fresh arenas, one per step (capped by memory: reuse round-robin beyond the cap)
    free = torch.cuda.mem_get_info()[0]
    cap = max(3, int(free * 0.6 // (n * 56)))
    R = min(K + W, cap)
    arenas = [eng.Population(desc, n).init(seed=1000 + r, slot0=rank * n) for r in range(R)]
    torch.cuda.synchronize()

    # ---- FP32 issue-rate probe = roofline denominator (lane-FFMA/s)
    probe = max(eng.fp32_probe(variant=0), eng.fp32_probe(variant=1))

    # ---- device-resident timing
    for s in range(W):
        arenas[s % R].self_apply(N_STEPS)
    barrier()
    sampler = ClockSampler(local_rank)
    if rank == 0:
        sampler.start()
    evs = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(K)]
    t_begin, t_end = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    t_begin.record()
    for s in range(K):
        evs[s][0].record()
        arenas[(W + s) % R].self_apply(N_STEPS)
        evs[s][1].record()
    t_end.record()
    barrier()
    clocks = sampler.stop() if rank == 0 else None
    total_ms = t_begin.elapsed_time(t_end)
    kern_ms = float(np.mean([a.elapsed_time(b) for a, b in evs]))
    t = torch.tensor([total_ms, kern_ms], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    total_ms, kern_ms = t.tolist()
    value = world * n * N_STEPS * K / (total_ms * 1e-3)

    # ---- end to end through the host API with HOST buffers, every step: pinned host records -> device (H2D stream),
    #      kernel + fused classification (compute stream), final weights AND class counters -> pinned host (D2H stream);
    #      three arenas in flight so the two copy directions and the kernel overlap
    NB = 3
    del arenas[NB:]
    torch.cuda.empty_cache()
    host_in = torch.empty((n, 14), dtype=torch.float32).pin_memory()
    lim = np.array([1.0] * 8 + [1.5 ** .5] * 4 + [2 ** .5] * 2)
    host_in.copy_(torch.from_numpy((np.random.default_rng(rank).uniform(-1, 1, (n, 14)) * lim).astype(np.float32)))
    host_out = [torch.empty((n, 14), dtype=torch.float32).pin_memory() for _ in range(2)]
    cnt_host = torch.zeros((NB, 5), dtype=torch.int64).pin_memory()
    cnt_dev = [None] * NB
    h2d, comp, d2h = torch.cuda.Stream(), torch.cuda.Stream(), torch.cuda.Stream()
    copied = [torch.cuda.Event() for _ in range(NB)]
    done = [torch.cuda.Event() for _ in range(NB)]
    freed = [torch.cuda.Event() for _ in range(NB)]

    def e2e_steps(count):
        for s in range(count):
            b = s % NB
            with torch.cuda.stream(h2d):
                h2d.wait_event(freed[b])
                arenas[b].w.copy_(host_in, non_blocking=True)
                copied[b].record(h2d)
            with torch.cuda.stream(comp):
                comp.wait_event(copied[b])
                cnt_dev[b] = arenas[b].self_apply(N_STEPS, want_counters=True, eps=1e-4)["counters"]
                done[b].record(comp)
            with torch.cuda.stream(d2h):
                d2h.wait_event(done[b])
                host_out[s % 2].copy_(arenas[b].w, non_blocking=True)
                cnt_host[b].copy_(cnt_dev[b], non_blocking=True)
                freed[b].record(d2h)
        d2h.synchronize()

    for b in range(NB):
        freed[b].record(d2h)
    e2e_steps(W)
    barrier()
    t0 = time.perf_counter()
    e2e_steps(K)
    barrier()
    e2e_s = time.perf_counter() - t0
    t = torch.tensor([e2e_s], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    e2e_value = world * n * N_STEPS * K / t.item()
    e2e_classes = dict(zip(eng.CLS_NAMES, cnt_host[(K - 1) % NB].tolist()))
    del host_in, host_out
    del arenas[:]
    torch.cuda.empty_cache()

    extra = {} if args.no_legs else secondary_legs(eng, torch, dist, rank, world, dev, probe)

    if rank == 0:
        achieved_tf = FLOP_PER_SELF_APP * n * N_STEPS / (kern_ms * 1e-3) / 1e12
        peak_tf = probe * 2.0 / 1e12
        sms = torch.cuda.get_device_properties(dev).multi_processor_count
        theo_tf = sms * 128 * 2 * (clocks["sm_max_mhz"] or 1965.0) * 1e6 / 1e12
        line = {
            "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": K, "warmup": W,
            "ms_per_step": total_ms / K, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
            "dtype": "f32", "data": "synthetic", "config": headline_config(n),
            "run": {"fresh_arenas": R, "init_rng": "Philox4x32-10 glorot-uniform on the device",
                    "host_affinity": ("%s, %d cpus" % (host_policy, len(host_cpus))) if host_cpus else "unbound"},
            "clocks": clocks,
            "e2e": {"value": e2e_value, "unit": UNIT, "h2d_bytes_per_step": n * 56, "d2h_bytes_per_step": n * 56 + 40,
                    "returns": "final weights of every net + the 5 class counters", "classes_last_step": e2e_classes},
            "gpu_launches": K,
            "roofline": {"bound": "fp32", "achieved": achieved_tf, "peak": peak_tf, "unit": "TFLOP/s",
                         "frac": achieved_tf / peak_tf, "traffic": NCU_TRAFFIC_BYTES_PER_NET * n,
                         "traffic_source": "ncu --set full, profiles/r02_k1_self_apply_fixed_ncu_summary.csv "
                                           "(algorithmic bytes per launch = 112 B x nets)",
                         "kernel": "self_apply_fixed_kernel<WW(2,2)>", "kernel_ms": kern_ms,
                         "peak_source": "FFMA-only probe kernel measured live (MEASURED_PEAKS.json has no FP32 entry)",
                         "peak_theoretical": theo_tf, "frac_vs_theoretical": achieved_tf / theo_tf,
                         "peak_theoretical_source": "SMs x 128 lanes x 2 FLOP x max SM clock",
                         "fma_pipe_frac": FP32_INSTR_PER_SELF_APP * n * N_STEPS / (kern_ms * 1e-3) / probe,
                         "hbm_gbs": BYTES_PER_NET * n / (kern_ms * 1e-3) / 1e9},
        }
        line.update(extra)
        if world == 1:
            import oracle
            oracle.set_threads(host_cores())
            v, cores, sample = cpu_port_rate(1 << 22, args.cpu_seconds)
            from oracle import numpy_twin
            line["cpu_baseline"] = {"value": v, "unit": UNIT, "cores": cores, "kind": "port", "sample": sample,
                                    "reference_structured_1core": numpy_twin.self_applications_per_second(),
                                    "reference_structured_note": "per-object Python loops of code/network.py:239-279 with a "
                                    "NumPy dot instead of model.predict, one core; the reference's own Keras/TF1 path is "
                                    "unavailable on this box"}
        print(json.dumps(line), flush=True)
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
