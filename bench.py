#!/usr/bin/env python
"""bench.py — events/sec of the keyed window-aggregate + alert hot path on B200.

Contract (see the task statement):  python bench.py --gpus N --steps K --warmup W
prints ONE JSON line.  A "step" is one micro-batch of the hot path:
    fmq_push_batch(ts, key, val, FMQ_WM_AUTO)  =  ingest (pane assign + late drop + per-event
    filter + keyed accumulate) -> plan (watermark) -> fire (emit windows + alert) -> commit.
Workload (BASELINE.json configs[2]): 10 s tumbling event-time window,
count/sum(avg)/max per host, 1 M hosts, watermark bound 5 s with a late tail, alert avg > limit,
plus the chapter1 per-event `> 90` filter; f64 values.

  value : whole-job events/s, inputs already resident in HBM, CUDA events on the engine's stream
  e2e   : same through the C ABI with pinned HOST columns (H2D inside the timed region) and the
          step's fired rows / alerts read back to the host (D2H inside)
  roofline : the ingest kernel, algorithmic bytes (24 B/event + 24 B/event-alert row) over its
          measured launch time, against MEASURED_PEAKS.json hbm_gbs
  cpu_baseline : the multi-thread C++ restatement of the Flink semantics (oracle/, kind "port")
          on a bounded sample of the same workload, on this box's host cores

--impl reference times that same CPU restatement (the reference's Flink job cannot run here: no
JVM) with all host threads.
"""
import argparse
import json
import os
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

import numpy as np  # noqa: E402

N_KEYS = 1_000_000
WINDOW_MS = 10_000
BOUND_MS = 5_000
AVG_LIMIT = 55.0
EVENT_LIMIT = 90.0


def workload_kwargs(capi, batch, n_keys=N_KEYS):
    return dict(window_enabled=1, agg_mask=capi.FMQ_AGG_COUNT | capi.FMQ_AGG_SUM | capi.FMQ_AGG_MAX,
                window_size_ms=WINDOW_MS, window_slide_ms=WINDOW_MS, max_out_of_orderness_ms=BOUND_MS,
                walert_source=capi.FMQ_ALERT_AVG, walert_op=capi.FMQ_CMP_GT, walert_threshold=AVG_LIMIT,
                efilter_enabled=1, efilter_op=capi.FMQ_CMP_GT, efilter_threshold=EVENT_LIMIT,
                emit_fired=1, table_capacity=int(n_keys * 1.5), pane_ring=16, max_batch=batch)


def stream_spec(StreamSpec, n_keys=N_KEYS, world=1):
    # 1000 events/ms per GPU over 1 M hosts = one report per host per second per GPU; jitter <= bound, 0.1 % late tail.
    # Weak scaling: a streaming job adds GPUs to keep up with a higher event RATE, so the global stream carries
    # world x 1000 events/ms (same hosts reporting world x as often) and one step still spans ~16.8 s of event time.
    return StreamSpec(n_keys=n_keys, events_per_ms=1000 * world, jitter_ms=BOUND_MS, late_frac_ppm=1000, late_extra_ms=25_000,
                      value_type="f64", key_kind="uniform", seed=42)


def bind_host_to_gpu_numa(index):
    """Pin this process to the CPUs of the NUMA node its GPU hangs off, before any pinned host buffer is
    allocated (first touch puts the pages there).  At N=8 every rank streams ~50 GB/s out of host memory;
    pages on the other socket halve that.  Returns a description for the JSON line (None if unknown)."""
    try:
        import torch
        pr = torch.cuda.get_device_properties(index)
        bus = "%04x:%02x:%02x.0" % (pr.pci_domain_id, pr.pci_bus_id, pr.pci_device_id)
        node = int(open(f"/sys/bus/pci/devices/{bus}/numa_node").read())
        if node < 0:
            return None
        cpus = set()
        for part in open(f"/sys/devices/system/node/node{node}/cpulist").read().strip().split(","):
            lo, _, hi = part.partition("-")
            cpus.update(range(int(lo), int(hi or lo) + 1))
        cpus &= os.sched_getaffinity(0)
        if not cpus:
            return None
        os.sched_setaffinity(0, cpus)
        return {"numa_node": node, "cpus": len(cpus), "pci": bus}
    except Exception:
        return None


class ClockSampler:
    """SM clock and throttle reasons while the timed region runs, read through NVML in a thread.
    (An `nvidia-smi -lms` child process was measured to stall every synchronising CUDA call of the
    sharded step — 27 ms/step instead of 2.1 — so it is only the fallback.)"""
    REASONS = {0x8: "hw_slowdown", 0x40: "hw_thermal_slowdown", 0x20: "sw_thermal_slowdown", 0x4: "sw_power_cap"}

    def __init__(self, index, period_s=0.004):     # the timed region of the default run is ~14 ms
        self.samples, self.stop_flag, self.index, self.period = [], False, index, period_s
        self.nvml = None
        try:
            import pynvml
            pynvml.nvmlInit()
            self.nvml = pynvml
            # CUDA_VISIBLE_DEVICES-relative index -> NVML handle by UUID-free approach: same order unless remapped
            vis = os.environ.get("CUDA_VISIBLE_DEVICES")
            phys = int(vis.split(",")[index]) if vis and all(x.strip().isdigit() for x in vis.split(",")) else index
            self.h = pynvml.nvmlDeviceGetHandleByIndex(phys)
            self.max_mhz = float(pynvml.nvmlDeviceGetMaxClockInfo(self.h, pynvml.NVML_CLOCK_SM))
        except Exception:
            self.nvml = None
        self.t = threading.Thread(target=self._run, daemon=True)
        self.t.start()

    def _run(self):
        if self.nvml is None:
            return
        n = self.nvml
        while not self.stop_flag:
            try:
                mhz = float(n.nvmlDeviceGetClockInfo(self.h, n.NVML_CLOCK_SM))
                try:
                    rs = int(n.nvmlDeviceGetCurrentClocksEventReasons(self.h))
                except Exception:
                    rs = int(n.nvmlDeviceGetCurrentClocksThrottleReasons(self.h))
                self.samples.append((time.time(), mhz, rs))
            except Exception:
                pass
            time.sleep(self.period)

    def stop(self, t0, t1):
        self.stop_flag = True
        self.t.join(timeout=1.0)
        if self.nvml is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvml unavailable"], "samples": 0}
        inside = [(m, r) for t, m, r in self.samples if t0 - 0.02 <= t <= t1 + 0.02] or [(m, r) for _, m, r in self.samples]
        reasons = set()
        for _, r in inside:
            for bit, name in self.REASONS.items():
                if r & bit:
                    reasons.add(name)
        return {"sm_mhz": float(np.median([m for m, _ in inside])) if inside else None, "sm_max_mhz": self.max_mhz,
                "reasons": sorted(reasons), "samples": len(inside), "source": "nvml"}


def measured_peak():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        return float(json.load(open(p))["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs)"
    return 6650.0, "fallback (B200_PROFILING.md 6.65 TB/s)"


def ncu_traffic():
    """dram bytes per ingest launch from the committed ncu --set full summary, if any."""
    p = os.path.join(ROOT, "profiles", "ncu_ingest_latest.json")
    if os.path.exists(p):
        try:
            return json.load(open(p)).get("dram_bytes_per_launch")
        except Exception:
            return None
    return None


# --------------------------------------------------------------------------------------------------
def cpu_reference_run(steps, warmup, sample_events, threads, capi, ParallelOracle, StreamSpec, generate_numpy):
    """The CPU restatement on host cores: `steps` timed steps of `sample_events` events each."""
    kw = workload_kwargs(capi, sample_events)
    spec = stream_spec(StreamSpec)
    eng = ParallelOracle(threads=threads, **kw)
    batches = [generate_numpy(spec, b * sample_events, sample_events) for b in range(warmup + steps)]
    for b in range(warmup):
        eng.push_batch(*batches[b]); eng.discard_outputs()
    t0 = time.perf_counter()
    for b in range(warmup, warmup + steps):
        eng.push_batch(*batches[b]); eng.discard_outputs()
    dt = time.perf_counter() - t0
    c = eng.counters()
    return steps * sample_events / dt, dt, c


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=40)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--batch-log2", type=int, default=24, help="events per micro-batch per GPU (log2)")
    ap.add_argument("--cpu-sample-log2", type=int, default=23, help="events per step of the CPU baseline")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-e2e", action="store_true")
    ap.add_argument("--multi", default="combiner", choices=["combiner", "exchange", "alltoall", "peerpull"],
                    help="N>1: combiner = local accumulate + peer-memory fire (no event shuffle; dense ids); "
                         "exchange = keyBy exchange inside the library (scatter over NVLink + device flags); "
                         "alltoall / peerpull = round 1's host-driven exchanges (NCCL all-to-all / copies out of peer memory)")
    ap.add_argument("--no-verify", action="store_true", help="N>1: skip the comparison of the ranks' outputs with the CPU oracle")
    ap.add_argument("--no-variants", action="store_true")
    ap.add_argument("--no-clocks", action="store_true", help="experiment: do not sample nvidia-smi during the timed region")
    ap.add_argument("--keys", type=int, default=N_KEYS, help="experiment: number of distinct hosts")
    ap.add_argument("--load", type=float, default=1.5, help="experiment: directory slots per key")
    ap.add_argument("--key-mode", default="dense", choices=["hashed", "dense"],
                    help="hashed: arbitrary u64 keys through the hash directory; dense: dictionary ids index the table")
    ap.add_argument("--no-efilter", action="store_true", help="experiment: drop the per-event filter from the workload")
    args = ap.parse_args()
    args.warmup = max(args.warmup, 3) if args.impl == "b200" else args.warmup

    rank = int(os.environ.get("RANK", "0")); world = int(os.environ.get("WORLD_SIZE", "1")); local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    from monitor_systam_flink_quickstart_b200 import capi
    from monitor_systam_flink_quickstart_b200.stream import StreamSpec, generate_numpy, generate_torch

    # ---------------------------------------------------------------- reference arm (CPU) ---------
    if args.impl == "reference":
        if rank != 0:
            return
        import __graft_entry__ as ge
        subprocess.check_call(["make", "-s", "-C", os.path.join(ROOT, "oracle")])
        from oracle_lib import ParallelOracle
        threads = os.cpu_count() or 1
        sample = 1 << args.cpu_sample_log2
        v, dt, c = cpu_reference_run(args.steps, args.warmup, sample, threads, capi, ParallelOracle, StreamSpec, generate_numpy)
        line = {"impl": "reference", "metric": "events/sec keyed window-agg+alert", "value": v, "unit": "events/s", "n_gpus": args.gpus,
                "steps": args.steps, "warmup": args.warmup, "ms_per_step": dt / args.steps * 1e3, "higher_is_better": True, "scaling": "weak",
                "vs_baseline": None, "dtype": "f64", "data": "synthetic",
                "config": {"workload": "C3 10s tumbling event-time count/avg/max per host + late drop + avg>limit alert + per-event >90 filter, 1M keys",
                           "events_per_step": sample, "note": "CPU restatement of the Flink semantics (oracle/cpu_parallel.cpp); Flink itself cannot run here (no JVM)"},
                "cpu_baseline": {"value": v, "unit": "events/s", "cores": threads, "kind": "port", "sample": f"{args.steps} steps x 2^{args.cpu_sample_log2} events"},
                "e2e": {"value": v, "unit": "events/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}}
        print(json.dumps(line))
        return

    # ---------------------------------------------------------------- B200 arm --------------------
    import torch
    if not torch.cuda.is_available():
        raise SystemExit("bench.py needs a CUDA device: the hot path has no CPU fallback")
    torch.cuda.set_device(local_rank)
    numa = bind_host_to_gpu_numa(local_rank) if world > 1 else None
    dist = None
    if world > 1:
        import torch.distributed as dist
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        dist.init_process_group("nccl", device_id=torch.device("cuda", local_rank))
    import __graft_entry__ as ge
    if rank == 0:
        ge.build()
    if dist:
        dist.barrier()

    B = 1 << args.batch_log2
    K, W = args.steps, args.warmup
    spec = stream_spec(StreamSpec, args.keys, world)
    kw = workload_kwargs(capi, B * (2 if world > 1 else 1), args.keys)
    kw["table_capacity"] = int(args.keys * args.load)
    kw["device"] = local_rank
    # size the pane ring for the event time one global step spans (panes per step + bound + late tail), power of two
    panes_per_step = (B * world // spec.events_per_ms) // WINDOW_MS + 4
    ring = 16
    while ring < panes_per_step + 4:
        ring *= 2
    kw["pane_ring"] = ring
    if args.no_efilter:
        kw["efilter_enabled"] = 0
    if args.key_mode == "dense":
        kw["key_mode"] = capi.FMQ_KEYS_DENSE
        # all-to-all route: rank-local dense ids (id // world); combiner route: every rank holds the full id space
        kw["table_capacity"] = (args.keys + world - 1) // world if (world > 1 and args.multi != "combiner") else args.keys
    # outputs stay in HBM during the timed region (polled after it): size the buffers for K steps.
    # ~1 fired row per 10 events (one report per host per second, 10 s windows), ~10 % event alerts.
    per_rank_events = B * max(K, W)
    kw["fired_capacity"] = int(per_rank_events * 0.125) + 4 * N_KEYS
    if args.keys < N_KEYS:
        kw["fired_capacity"] = int(per_rank_events * 0.125 * 4) + 4 * N_KEYS
    kw["walert_capacity"] = kw["fired_capacity"]
    kw["ealert_capacity"] = int(per_rank_events * 0.115) + 65536
    # synthetic inputs, resident in HBM before any timing: rank r owns slice r of every global batch
    def dev_batch(step):
        start = (step * world + rank) * B
        ts, key, val = generate_torch(spec, start, B, device=f"cuda:{local_rank}")
        return ts.contiguous(), key.contiguous(), val.contiguous()
    default_batches = batches = [dev_batch(s) for s in range(W + K)]
    torch.cuda.synchronize()

    host_issue_ms = [0.0]

    def make_operator(eng, kw, multi):
        from monitor_systam_flink_quickstart_b200.shard import CombinerOperator, ExchangeOperator, PeerPullOperator, ShardedOperator
        if multi == "combiner" and kw.get("key_mode") == capi.FMQ_KEYS_DENSE:
            return CombinerOperator(eng, rank, world)
        if multi in ("exchange", "combiner"):
            return ExchangeOperator(eng, rank, world, B)
        if multi == "peerpull":
            return PeerPullOperator(eng, rank, world, BOUND_MS)
        return ShardedOperator(eng, rank, world, BOUND_MS)

    def measure(kw, with_clocks, multi=args.multi, batches=None, K=K):
        """W warm-up + K timed steps with device-resident inputs; CUDA events on the engine's stream
        (every kernel, copy and collective of a step is ordered on it)."""
        batches = batches if batches is not None else default_batches
        eng = capi.Engine(**kw)
        stream = torch.cuda.ExternalStream(eng.stream, device=torch.device("cuda", local_rank))
        sharded = make_operator(eng, kw, multi) if world > 1 else None

        def step(b):
            if sharded:
                sharded.process_batch(*batches[b])
            else:
                eng.push_batch(*batches[b])

        def drain():
            eng.synchronize()
            return len(eng.poll_fired()), len(eng.poll_window_alerts()), len(eng.poll_event_alerts())

        # the sampler thread starts before the warm-up so that the timed region follows the last
        # warm-up step without an idle gap (an idle gap of 0.25 s was measured to cost the first
        # NCCL step after it ~250 ms)
        sampler = ClockSampler(local_rank) if (with_clocks and rank == 0 and not args.no_clocks) else None
        for b in range(W):
            step(b)
        drain()
        c0 = eng.counters()
        if dist:
            dist.barrier()
        torch.cuda.synchronize()
        ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        t_wall0 = time.time()
        ev0.record(stream)
        t_issue0 = time.perf_counter()
        for b in range(W, W + K):
            step(b)
        host_issue_ms[0] = (time.perf_counter() - t_issue0) * 1e3 / K     # how long the host needs to queue one step
        if hasattr(sharded, "join"):
            sharded.join()                     # the last fire / retire run on a second stream: the clock stops after them
        ev1.record(stream)
        eng.synchronize()
        torch.cuda.synchronize()
        t_wall1 = time.time()
        if dist:
            dist.barrier()
        ms = ev0.elapsed_time(ev1)
        clocks = sampler.stop(t_wall0, t_wall1) if sampler else None
        # per-kernel timing over the timed pushes (CUDA events on the engine stream, recorded by the library)
        ing = [eng.push_timing(back)["ingest_ms"] for back in range(min(K, 64))]
        fire = [eng.push_timing(back)["fire_ms"] for back in range(min(K, 64))]
        c1 = eng.counters()
        delta = {k: c1[k] - c0[k] for k in ("gpu_launches", "event_alerts", "late_dropped", "windows_fired", "events_in")}
        drain()
        eng.close()
        return ms, float(np.mean(ing)), float(np.mean(fire)), delta, clocks

    ms, ing_ms, fire_ms, delta, clocks = measure(kw, True)
    issue_ms = host_issue_ms[0]
    launches, n_ealert, n_late, n_fired = delta["gpu_launches"], delta["event_alerts"], delta["late_dropped"], delta["windows_fired"]
    ealert_rank, events_rank = delta["event_alerts"] / K, delta["events_in"] / K
    if dist:
        t = torch.tensor([ms], device=f"cuda:{local_rank}", dtype=torch.float64)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        ms = float(t.item())
        tot = torch.tensor([launches, n_ealert, n_late, n_fired], device=f"cuda:{local_rank}", dtype=torch.int64)
        dist.all_reduce(tot)
        launches, n_ealert, n_late, n_fired = [int(x) for x in tot.tolist()]
    value = world * B * K / (ms * 1e-3)
    sharded = world > 1

    # the other key mode, same stream, same steps (N = 1 only): reported as a variant, not the headline
    variants = None
    if world == 1 and not args.no_variants:
        other = "hashed" if args.key_mode == "dense" else "dense"
        kw2 = dict(kw)
        kw2["key_mode"] = capi.FMQ_KEYS_DENSE if other == "dense" else capi.FMQ_KEYS_HASHED
        kw2["table_capacity"] = args.keys if other == "dense" else int(args.keys * args.load)
        ms2, ing2, fire2, d2, _ = measure(kw2, False)
        variants = {f"key_mode={other}": {"value": B * K / (ms2 * 1e-3), "unit": "events/s", "ms_per_step": ms2 / K, "ingest_kernel_ms": ing2,
                                          "roofline_frac": (24.0 * B + 24.0 * d2["event_alerts"] / K) / (ing2 * 1e-3) / 1e9 / measured_peak()[0]}}

        # the same stream through two narrower jobs (both inside the headline kernel's code path):
        #   C1 (BASELINE.json configs[0]): chapter1's per-event `> 90` filter alone — no table, the streaming skeleton itself
        #   count+sum: chapter3's own aggregate (BandwidthMonitor.java:37) — ONE accumulator request per event instead of two
        def narrow(tag, kwn, alg_rows):
            msn, ingn, firen, dn, _ = measure(kwn, False)
            bytes_ = 24.0 * B + 24.0 * dn["event_alerts"] / K
            variants[tag] = {"value": B * K / (msn * 1e-3), "unit": "events/s", "ms_per_step": msn / K, "ingest_kernel_ms": ingn,
                             "roofline_frac": bytes_ / (ingn * 1e-3) / 1e9 / measured_peak()[0], "algorithmic_bytes_per_launch": bytes_}
        kw_c1 = dict(kw); kw_c1.update(window_enabled=0, walert_source=0)
        narrow("workload=C1 per-event >90 filter only (BASELINE.json configs[0])", kw_c1, 0)
        kw_cs = dict(kw); kw_cs.update(agg_mask=capi.FMQ_AGG_COUNT | capi.FMQ_AGG_SUM, efilter_enabled=0)
        narrow("aggregate=count+sum only, no per-event filter (one RED request per event)", kw_cs, 0)

        # BASELINE.json configs[1] (chapter2 keyed ValueState rule, 100 K hosts) beside the headline:
        # resolve -> radix sort by slot -> segmented scan of the per-record transition functions
        spec_c2 = StreamSpec(n_keys=100_000)
        eng_c2 = capi.Engine(max_batch=B, consec_n=3, consec_threshold=90.0, consec_reset_on_fire=1, table_capacity=100_000,
                             key_mode=capi.FMQ_KEYS_DENSE, calert_capacity=B // 8, device=local_rank)
        t_c2 = []
        for i in range(3 + min(K, 5)):
            cols = generate_torch(spec_c2, i * B, B, device=f"cuda:{local_rank}")
            torch.cuda.synchronize()
            eng_c2.push_batch(*cols); eng_c2.synchronize()
            if i >= 3:
                t_c2.append(eng_c2.push_timing(0)["total_ms"])
            eng_c2.poll_consec_alerts(chunk=1 << 24)
        c2_launches = eng_c2.counters()["gpu_launches"]
        eng_c2.close()
        variants["workload=C2 3 consecutive readings > 90 per host, 100K hosts (BASELINE.json configs[1])"] = {
            "value": B / (float(np.mean(t_c2)) * 1e-3), "unit": "events/s", "ms_per_step": float(np.mean(t_c2)), "gpu_launches": int(c2_launches),
            "timing": "CUDA events on the engine stream around each push (library-recorded), inputs resident in HBM"}

    def job_max(x):
        if not dist:
            return x
        t = torch.tensor([x], device=f"cuda:{local_rank}", dtype=torch.float64)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item())

    verify = None
    if world > 1 and not args.no_variants:
        variants = {}
        nv_bytes = 24.0 * B * (world - 1) / world            # what one GPU sends (and receives) over NVLink per step
        # (a) the headline workload through the reference's own structure: a keyBy exchange in front of owner-held state
        kwx = dict(kw)
        if args.key_mode == "dense":
            kwx["table_capacity"] = (args.keys + world - 1) // world
        msx, ingx, firex, dx, _ = measure(kwx, False, multi="exchange")
        msx = job_max(msx)
        variants["route=keyBy exchange inside the library (scatter over NVLink + device flags), same C3 workload"] = {
            "value": world * B * K / (msx * 1e-3), "unit": "events/s", "ms_per_step": msx / K,
            "nvlink_GBps_per_gpu_each_direction": nv_bytes / (msx / K * 1e-3) / 1e9, "nvlink_frac_of_900": nv_bytes / (msx / K * 1e-3) / 1e9 / 900.0,
            "events_over_nvlink_per_step_per_gpu": B * (world - 1) // world}
        # (b) BASELINE.json configs[3] in its named form: 60 s / 10 s sliding sum per (host << 32 | metric), 10 M keys, int64, hashed keys
        K4 = min(K, 8)
        spec4 = StreamSpec(n_keys=10_000_000, events_per_ms=10_000 * world, jitter_ms=4000, late_frac_ppm=0, value_type="i64", key_kind="host_metric")
        keys_rank = 10_000_000 // world
        kw4 = dict(value_type=capi.FMQ_VAL_I64, window_enabled=1, agg_mask=15, window_size_ms=60_000, window_slide_ms=10_000, max_out_of_orderness_ms=5000,
                   walert_source=capi.FMQ_ALERT_SUM, walert_op=capi.FMQ_CMP_LT, walert_mul=8.0, walert_div0=60.0, walert_div1=1024.0, walert_div2=1024.0,
                   walert_threshold=0.05, table_capacity=int(keys_rank * 1.6) + 1024, pane_ring=16, max_batch=2 * B, device=local_rank,
                   fired_capacity=8 * keys_rank + (1 << 20), walert_capacity=8 * keys_rank + (1 << 20))
        batches4 = [tuple(t.contiguous() for t in generate_torch(spec4, (s_ * world + rank) * B, B, device=f"cuda:{local_rank}")) for s_ in range(W + K4)]
        torch.cuda.synchronize()
        ms4, ing4, fire4, d4, _ = measure(kw4, False, multi="exchange", batches=batches4, K=K4)
        ms4 = job_max(ms4)
        del batches4
        variants["workload=configs[3]: 60s/10s sliding per (host,metric), 10M keys, hashed, through the keyBy exchange"] = {
            "value": world * B * K4 / (ms4 * 1e-3), "unit": "events/s", "ms_per_step": ms4 / K4, "steps": K4,
            "nvlink_GBps_per_gpu_each_direction": nv_bytes / (ms4 / K4 * 1e-3) / 1e9, "nvlink_frac_of_900": nv_bytes / (ms4 / K4 * 1e-3) / 1e9 / 900.0}

        # (c) BASELINE.json configs[4]: Zipf(1.2) over 10^8 ids (one id carries 18 % of the stream, so does its owner rank),
        #     tumbling windows plus the consecutive-readings rule (order-dependent keyed state: stable exchange), hashed keys
        spec5 = StreamSpec(n_keys=100_000_000, events_per_ms=1000 * world, jitter_ms=3000, late_frac_ppm=500, late_extra_ms=20_000, key_kind="zipf")
        kw5 = dict(window_enabled=1, agg_mask=15, window_size_ms=10_000, window_slide_ms=10_000, max_out_of_orderness_ms=3000,
                   consec_n=3, consec_threshold=90.0, consec_reset_on_fire=1, table_capacity=6_000_000, pane_ring=16, max_batch=3 * B, device=local_rank,
                   fired_capacity=24_000_000, calert_capacity=3 * B)
        batches5 = [tuple(t.contiguous() for t in generate_torch(spec5, (s_ * world + rank) * B, B, device=f"cuda:{local_rank}")) for s_ in range(W + K4)]
        torch.cuda.synchronize()
        ms5, ing5, fire5, d5, _ = measure(kw5, False, multi="exchange", batches=batches5, K=K4)
        ms5 = job_max(ms5)
        del batches5
        variants["workload=configs[4]: Zipf(1.2) over 1e8 ids, tumbling window + 3 consecutive readings > 90, hashed, through the keyBy exchange"] = {
            "value": world * B * K4 / (ms5 * 1e-3), "unit": "events/s", "ms_per_step": ms5 / K4, "steps": K4,
            "note": "the hottest id's owner receives ~18 % of all events on top of its share: the step time is that rank's"}

    if world > 1 and not args.no_verify:
        # Outside every timed region: V steps of the headline route into fresh engines; the union of the ranks' rows
        # against the CPU oracle fed the same global batches, as order-independent sums over the rows.
        V = 2
        engv = capi.Engine(**kw)
        opv = make_operator(engv, kw, args.multi)
        for b in range(V):
            opv.process_batch(*default_batches[b])
        engv.synchronize()
        opv.flush()
        engv.synchronize()

        def row_sums(f, a, e, late, gk):
            M = np.uint64(0x9E3779B97F4A7C15)
            hf = ((gk(f.key) * M) ^ (f.window_end.view(np.uint64) * np.uint64(0xC2B2AE3D27D4EB4F)) ^ f.count.view(np.uint64)).sum(dtype=np.uint64)
            he = ((e.ts.view(np.uint64) * M) ^ gk(e.key) ^ e.val.view(np.uint64)).sum(dtype=np.uint64)
            s64 = lambda u: int(np.array(u, dtype=np.uint64).astype(np.int64))  # noqa: E731
            return [len(f), int(f.count.sum()), s64(hf), len(a), len(e), s64(he), int(late)]
        gk = opv.global_keys if hasattr(opv, "global_keys") else (lambda k: k)
        mine = row_sums(engv.poll_fired(chunk=1 << 24), engv.poll_window_alerts(chunk=1 << 24), engv.poll_event_alerts(chunk=1 << 24),
                        engv.counters()["late_dropped"], gk)
        engv.close()
        tot = torch.tensor(mine, device=f"cuda:{local_rank}", dtype=torch.int64)
        dist.all_reduce(tot)                                   # int64 sums wrap like the uint64 sums they stand for
        if rank == 0:
            subprocess.check_call(["make", "-s", "-C", os.path.join(ROOT, "oracle")])
            from oracle_lib import ParallelOracle
            okw = {k_: v_ for k_, v_ in kw.items() if k_ != "device"}
            orc = ParallelOracle(threads=os.cpu_count() or 1, **okw)
            for b in range(V):
                parts = [generate_numpy(spec, (b * world + r) * B, B) for r in range(world)]
                orc.push_batch(*[np.concatenate([p_[j] for p_ in parts]) for j in range(3)])
                del parts
            orc.flush()
            want = row_sums(orc.poll_fired(chunk=1 << 26), orc.poll_window_alerts(chunk=1 << 26), orc.poll_event_alerts(chunk=1 << 26),
                            orc.counters()["late_dropped"], lambda k: k)
            orc.close()
            got = [int(x) for x in tot.tolist()]
            names = ["fired_rows", "sum_of_counts", "fired_row_hash", "window_alert_rows", "event_alert_rows", "event_alert_hash", "late_dropped"]
            verify = {"steps": V, "events": V * world * B, "route": args.multi, "oracle": "oracle/cpu_parallel.cpp (ParallelOracle)",
                      "equal": got == want, "got": dict(zip(names, got)), "want": dict(zip(names, want))}

    # -------- roofline of the dominant kernel (ingest), this rank --------
    peak, peak_src = measured_peak()
    alg_bytes = 24.0 * events_rank + 24.0 * ealert_rank
    achieved = alg_bytes / (ing_ms * 1e-3) / 1e9
    roofline = {"bound": "hbm", "kernel": "ingest_ring_kernel<double,%s>" % args.key_mode, "achieved": achieved, "peak": peak, "unit": "GB/s", "frac": achieved / peak,
                "traffic": ncu_traffic(), "peak_source": peak_src, "algorithmic_bytes_per_launch": alg_bytes,
                "avg_launch_ms": ing_ms, "fire_chain_ms": fire_ms, "share_of_step": ing_ms * K / ms,
                "host_issue_ms_per_step": issue_ms}
    if world > 1 and args.multi == "combiner" and args.key_mode == "dense":
        roofline["fire_chain_note"] = "runs on a second stream beside the next batch's ingest: latency from the end of ingest to the end of retire, not part of the step's critical path"

    # -------- e2e: host columns through the C ABI, outputs read back --------
    # What a JVM shim would do per micro-batch: hand over three pinned host columns (direct
    # ByteBuffers), then drain every output column into pinned host buffers.  H2D and D2H copies are
    # inside the timed region; host wall clock with a full synchronise at both ends.
    e2e = None
    e2e_packed = None
    combiner = sharded and args.multi == "combiner" and args.key_mode == "dense"
    if not args.no_e2e and (not sharded or combiner):
        e2e_steps = min(K, 8)
        cap_rows = int(B * 0.14) + 2 * N_KEYS

        def pinned(n, dtype):
            return torch.empty(n, dtype=dtype).pin_memory()

        def to_host(cols, packed):
            ts, key, val = (t.cpu() for t in cols)
            if not packed:
                return (ts.pin_memory(), key.pin_memory(), val.pin_memory(), 0)
            # the narrow wire format of fmq_push_packed: uint32 ms offsets from the batch's first
            # timestamp, uint32 host ids, the double as it is: 16 B per event instead of 24
            base = int(ts.min())
            assert int(ts.max()) - base < (1 << 32) and int(key.max()) < (1 << 32)
            ts32 = torch.from_numpy((ts.numpy() - base).astype(np.uint32).view(np.int32)).pin_memory()
            key32 = torch.from_numpy(key.numpy().astype(np.uint32).view(np.int32)).pin_memory()
            return (ts32, key32, val.pin_memory(), base)

        def run_e2e(packed):
            host = [to_host(batches[W + s], packed) for s in range(e2e_steps)]
            warm = [to_host(batches[s], packed) for s in range(W)]
            eng2 = capi.Engine(overlapped_polls=1, **kw)
            op2 = None
            if combiner:
                from monitor_systam_flink_quickstart_b200.shard import CombinerOperator
                op2 = CombinerOperator(eng2, rank, world)
            # two sets of pinned output columns: the copy-out of step s runs while step s+1 is copied in
            out_sets = [([pinned(cap_rows, torch.int64) for _ in range(3)] + [pinned(cap_rows, torch.float64) for _ in range(3)],
                         [pinned(cap_rows, torch.int64), pinned(cap_rows, torch.int64), pinned(cap_rows, torch.float64)],
                         [pinned(cap_rows, torch.int64), pinned(cap_rows, torch.int64), pinned(cap_rows, torch.float64)]) for _ in range(2)]
            step_no = [0]

            def np_cols(cols):
                ts, key, val, base = cols
                if packed:
                    return (ts.numpy().view(np.uint32), key.numpy().view(np.uint32), val.numpy(), base)
                return (ts.numpy(), key.numpy().view(np.uint64), val.numpy(), 0)

            def e2e_step(cols, nxt):
                ts, key, val, base = cols
                if op2:                                                                     # own events in, 2 x all-reduce, fire over NVLink
                    op2.process_batch(ts, key, val, ts_base=base)
                else:
                    eng2.push_packed(ts, key, val, ts_base=base)                           # H2D (already under way); returns when the copy is done
                if nxt is not None:
                    eng2.stage_host(*nxt[:3], ts_base=nxt[3])                              # a later batch's H2D is queued: the copy engine never waits for the host
                eng2.poll_all_end()                                                         # previous step's rows are now on the host
                nf, na, ne = eng2.poll_all_begin(*out_sets[step_no[0] % 2])                # waits for this step's kernels, starts its D2H
                step_no[0] += 1
                return 48 * nf + 24 * na + 24 * ne
            seq = [np_cols(c) for c in warm + host]
            for i in range(W):
                e2e_step(seq[i], None)
            eng2.poll_all_end()
            d2h = 0
            torch.cuda.synchronize()
            if dist:
                dist.barrier()
            t0 = time.perf_counter()
            eng2.stage_host(*seq[W][:3], ts_base=seq[W][3])                                 # inside the timed region: two batches are staged ahead
            if e2e_steps > 1:
                eng2.stage_host(*seq[W + 1][:3], ts_base=seq[W + 1][3])
            for s in range(e2e_steps):
                d2h += e2e_step(seq[W + s], seq[W + s + 2] if s + 2 < e2e_steps else None)
            eng2.poll_all_end()
            eng2.synchronize()
            dt = time.perf_counter() - t0
            if dist:                                                                        # whole job: slowest rank's time, every rank's rows
                t = torch.tensor([dt, -dt], device=f"cuda:{local_rank}", dtype=torch.float64)
                dist.all_reduce(t, op=dist.ReduceOp.MAX)
                dt = float(t[0].item())
                tot = torch.tensor([d2h], device=f"cuda:{local_rank}", dtype=torch.int64)
                dist.all_reduce(tot)
                d2h = int(tot.item())
            eng2.close()
            per_event = sum(c.element_size() for c in host[0][:3])
            return {"value": e2e_steps * B * world / dt, "unit": "events/s", "h2d_bytes_per_step": per_event * B * world, "d2h_bytes_per_step": d2h // e2e_steps,
                    "steps": e2e_steps, "ms_per_step": dt / e2e_steps * 1e3, "host_numa_binding": numa,
                    "entry_point": ("CombinerOperator: fmq_push_local_packed + fmq_stage_host + 2 all-reduces + fmq_fire_global/fmq_retire, " if combiner else "fmq_push_packed + fmq_stage_host, ") + ("uint32 ts offset + uint32 host id + double" if packed else "int64 ts + uint64 key + double (the fmq_push_batch format)"),
                    "timing": "host wall clock around K x [push of pinned host columns + fmq_poll_all_begin/end into pinned host buffers]; "
                              "the D2H of step s overlaps the H2D of step s+1; every row is on the host when the clock stops"}
        # What a JVM shim would do per micro-batch: hand over three pinned host columns (direct
        # ByteBuffers), then drain every output column into pinned host buffers.  H2D and D2H copies are
        # inside the timed region; host wall clock with a full synchronise at both ends.  The host path
        # headline = the 24 B/event columns of the boundary (int64 ts, uint64 key, 8-byte value);
        # the 16 B/event packed wire format is reported beside it
        e2e = run_e2e(False)
        e2e_packed = run_e2e(True)

    if rank != 0:
        if dist:
            dist.destroy_process_group()
        return

    # -------- CPU baseline on a bounded sample (rank 0, N = 1 only) --------
    cpu_baseline = None
    if not args.no_cpu_baseline and world == 1:
        subprocess.check_call(["make", "-s", "-C", os.path.join(ROOT, "oracle")])
        from oracle_lib import ParallelOracle
        threads = os.cpu_count() or 1
        v, dt, _ = cpu_reference_run(12, 1, 1 << args.cpu_sample_log2, threads, capi, ParallelOracle, StreamSpec, generate_numpy)
        # the single-thread literal restatement (oracle/flink_literal.cpp: ordered map of (key, window), explicit
        # timers) on a smaller sample: what one core does per-record, beside the key-partitioned port above
        from oracle_lib import OracleEngine
        lit_n = 1 << 21
        lit = OracleEngine(**workload_kwargs(capi, lit_n))
        lit_batches = [generate_numpy(stream_spec(StreamSpec), b * lit_n, lit_n) for b in range(3)]
        lit.push_batch(*lit_batches[0])
        t_l = time.perf_counter()
        for b in (1, 2):
            lit.push_batch(*lit_batches[b])
        lit_v = 2 * lit_n / (time.perf_counter() - t_l)
        lit.close()
        cpu_baseline = {"value": v, "unit": "events/s", "cores": threads, "kind": "port",
                        "single_thread_literal_oracle": {"value": lit_v, "unit": "events/s", "cores": 1, "sample": "2 steps x 2^21 events after 1 warm-up step"},
                        "sample": f"12 steps x 2^{args.cpu_sample_log2} events of the same stream: {dt:.1f} s wall x {threads} threads = {dt * threads:.0f} core-seconds"}

    line = {"metric": "events/sec keyed window-agg+alert", "value": value, "unit": "events/s", "n_gpus": world, "steps": K, "warmup": W,
            "ms_per_step": ms / K, "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f64", "data": "synthetic",
            "config": {"workload": "C3 10s tumbling event-time count/avg/max per host + late drop + avg>limit alert + per-event >90 filter, 1M keys",
                       "events_per_step_per_gpu": B, "global_events_per_ms": spec.events_per_ms, "keys": args.keys, "key_mode": args.key_mode, "window_ms": WINDOW_MS, "bound_ms": BOUND_MS,
                       "l2": "inputs larger than L2 (each batch = 24 B x 2^%d = %.0f MB, read once)" % (args.batch_log2, 24 * B / 1e6),
                       "parallelism": "1 GPU" if world == 1 else (
                           f"{world} GPUs, combiner: local accumulate + 32 B of batch scalars exchanged by peer stores behind device flags + fire over NVLink peer memory, fire/retire of batch k overlapped with the ingest of batch k+1"
                           if (args.multi == "combiner" and args.key_mode == "dense") else (
                               f"keyBy exchange over {world} GPUs inside the library: owner runs scattered over NVLink, device flags" if args.multi == "exchange" else f"keyBy sharded over {world} GPUs, buckets pulled from peer memory over NVLink" if args.multi == "peerpull"
                               else f"keyBy hash-sharded over {world} GPUs, NCCL all-to-all"))},
            "roofline": roofline, "cpu_baseline": cpu_baseline, "e2e": e2e, "e2e_packed": e2e_packed, "gpu_launches": int(launches), "clocks": clocks, "variants": variants, "verify_vs_oracle": verify,
            "step_counts": {"event_alert_rows": int(n_ealert), "late_dropped": int(n_late), "fired_rows": int(n_fired)}}
    print(json.dumps(line))
    if dist:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
