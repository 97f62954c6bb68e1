#!/usr/bin/env python
"""Headline benchmark: env-steps/s of the batched MyoHand physics step (BASELINE.json metric).

  python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference]
  python -m torch.distributed.run --nnodes=1 --nproc-per-node N ... bench.py --gpus N ...

One "step" = one mj_step of every world of this rank (one launch of the fused kernel), on the MyoHand model of
myosuite/mjx/play.py:8 with 8192 worlds per GPU, random ctrl ~ U[0,1] redrawn every 10 steps.
Prints ONE JSON line (rank 0).  `--impl reference` times the CPU restatement (oracle/) on the host cores instead:
the reference's own CPU path (MuJoCo C through dm_control) is third-party code that is neither vendored nor
installed here, so the "port" is the only CPU arm that exists.
"""
import argparse
import json
import os
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

FRAME_SKIP = 10
SEED = 1234


def load_model(name):
    from myosuite_mjx_b200 import mjcf
    return mjcf.Model.load(os.path.join(ROOT, "tests", "golden", "models", name + ".npz"))


def initial_qpos(m, nworld, seed):
    """qpos0 + U(-0.1, 0.1) * range, clipped to the joint range."""
    rng = np.random.default_rng(seed)
    q = np.tile(np.asarray(m.qpos0, dtype=np.float64), (nworld, 1))
    for j in range(m.njnt):
        if m.jnt_type[j] in (2, 3) and m.jnt_limited[j]:
            a = m.jnt_qposadr[j]
            lo, hi = m.jnt_range[j]
            q[:, a] = np.clip(q[:, a] + rng.uniform(-0.1, 0.1, nworld) * (hi - lo), lo, hi)
    return q


class ClockSampler(threading.Thread):
    """Samples SM clock and throttle reasons through NVML while the timed region runs."""

    def __init__(self, index):
        super().__init__(daemon=True)
        self.index, self.samples, self.reasons, self.max_mhz, self._stop_evt = index, [], set(), None, threading.Event()
        try:
            import pynvml
            pynvml.nvmlInit()
            self.nv = pynvml
            self.h = pynvml.nvmlDeviceGetHandleByIndex(index)
            self.max_mhz = pynvml.nvmlDeviceGetMaxClockInfo(self.h, pynvml.NVML_CLOCK_SM)
        except Exception:
            self.nv = None

    def run(self):
        if self.nv is None:
            return
        nv = self.nv
        names = {nv.nvmlClocksThrottleReasonSwPowerCap: "sw_power_cap", nv.nvmlClocksThrottleReasonHwSlowdown: "hw_slowdown",
                 nv.nvmlClocksThrottleReasonHwThermalSlowdown: "hw_thermal_slowdown",
                 nv.nvmlClocksThrottleReasonSwThermalSlowdown: "sw_thermal_slowdown",
                 nv.nvmlClocksThrottleReasonHwPowerBrakeSlowdown: "hw_power_brake"}
        while not self._stop_evt.is_set():
            try:
                self.samples.append(nv.nvmlDeviceGetClockInfo(self.h, nv.NVML_CLOCK_SM))
                r = nv.nvmlDeviceGetCurrentClocksThrottleReasons(self.h)
                for bit, nm in names.items():
                    if r & bit:
                        self.reasons.add(nm)
            except Exception:
                pass
            time.sleep(0.02)

    def stop(self):
        self._stop_evt.set()
        self.join(timeout=1)
        return {"sm_mhz": float(np.median(self.samples)) if self.samples else None, "sm_max_mhz": self.max_mhz,
                "reasons": sorted(self.reasons)}


def cpu_baseline(m, cores, target_seconds=12.0):
    """Oracle port on the host cores over a bounded sample of the same workload."""
    sys.path.insert(0, os.path.join(ROOT, "oracle"))
    import oracle
    nworld = 4 * cores
    q0 = initial_qpos(m, nworld, SEED)
    t = time.time()
    oracle.rollout(m, nworld, 50, FRAME_SKIP, q0, SEED, cores)      # calibrate
    rate = nworld * 50 / max(time.time() - t, 1e-6)
    nsteps = int(max(50, min(2000, target_seconds * rate / nworld)))
    t = time.time()
    oracle.rollout(m, nworld, nsteps, FRAME_SKIP, q0, SEED, cores)
    dt = time.time() - t
    return {"value": nworld * nsteps / dt, "unit": "env-steps/s", "cores": cores, "kind": "port",
            "sample": f"{nworld} worlds x {nsteps} mj_steps, random ctrl held {FRAME_SKIP} steps, float64, {cores} threads"}


def run_reference(args, rank, world):
    """CPU arm: the oracle port with every host thread (rank 0 only)."""
    if rank != 0:
        return
    m = load_model(args.model)
    cores = os.cpu_count() or 1
    sys.path.insert(0, os.path.join(ROOT, "oracle"))
    import oracle
    nworld = 4 * cores
    per_step = 20   # mj_steps per bench "step" for the bounded CPU sample
    q0 = initial_qpos(m, nworld, SEED)
    for _ in range(args.warmup):
        oracle.rollout(m, nworld, per_step, FRAME_SKIP, q0, SEED, cores)
    t = time.time()
    for _ in range(args.steps):
        oracle.rollout(m, nworld, per_step, FRAME_SKIP, q0, SEED, cores)
    dt = time.time() - t
    val = nworld * per_step * args.steps / dt
    sample = f"{nworld} worlds x {per_step} mj_steps per step, {cores} threads, float64 CPU restatement (oracle/), not MuJoCo"
    line = {"impl": "reference", "metric": "env-steps/s", "value": val, "unit": "env-steps/s", "n_gpus": args.gpus, "steps": args.steps,
            "warmup": args.warmup, "ms_per_step": 1e3 * dt / args.steps, "higher_is_better": True, "scaling": "weak",
            "vs_baseline": None, "dtype": "f64", "data": "synthetic",
            "config": {"workload": WORKLOAD.format(model=args.model, nworld=args.nworld)},
            "cpu_baseline": {"value": val, "unit": "env-steps/s", "cores": cores, "kind": "port", "sample": sample},
            "e2e": {"value": val, "unit": "env-steps/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}}
    print(json.dumps(line), flush=True)


WORKLOAD = "MyoHand ({model}.xml, play.py:8) random-ctrl rollout, {nworld} worlds/GPU, dt 0.002, Newton, full mjData written every step"


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=1000)
    ap.add_argument("--warmup", type=int, default=100)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--model", default="hand")
    ap.add_argument("--nworld", type=int, default=8192, help="worlds per GPU")
    ap.add_argument("--fields", default="all", choices=["all", "core"], help="all: drop-in mjx.step (full Data); core: state only")
    ap.add_argument("--nsub", type=int, default=1, help="mj_steps fused per launch")
    ap.add_argument("--no-cpu", action="store_true", help="skip the cpu_baseline leg")
    ap.add_argument("--pipelined", action="store_true", help="opt-in: wide-output kernel of step k under the kernel of step k+1 (myo_batch_set_pipelined)")
    args = ap.parse_args()
    rank = int(os.environ.get("RANK", "0")); world = int(os.environ.get("WORLD_SIZE", "1")); local = int(os.environ.get("LOCAL_RANK", "0"))
    if args.impl == "reference":
        run_reference(args, rank, world)
        return

    import torch
    import torch.distributed as dist
    from myosuite_mjx_b200 import lib, mjx
    if not torch.cuda.is_available():
        raise SystemExit("bench.py needs a CUDA device: the product has no CPU fallback")
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)
    m = load_model(args.model)
    model = mjx.put_model(m, device=local)
    nw = args.nworld
    # worlds are partitioned by index: rank r owns [r*nw, (r+1)*nw); seeds depend on the global world index only
    q0 = initial_qpos(m, nw * world, SEED)[rank * nw:(rank + 1) * nw]
    d = mjx.put_data(model, dict(qpos=q0), nworld=nw, fields=args.fields)
    if args.pipelined:
        d.set_pipelined(True)
    ctrl_buf = d.storage("ctrl")
    gen = torch.Generator(device=dev); gen.manual_seed(SEED + rank)
    stream = torch.cuda.current_stream()

    def one_step(i):
        if i % FRAME_SKIP == 0:
            ctrl_buf.uniform_(0.0, 1.0, generator=gen)
        mjx.step(model, d, args.nsub)

    for i in range(args.warmup):
        one_step(i)
    torch.cuda.synchronize()
    if world > 1:
        dist.barrier()
    torch.cuda.synchronize()
    sampler = ClockSampler(local); sampler.start()
    launches0 = lib.lib().myo_launch_count()
    ev = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(args.steps)]
    t_all0, t_all1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    t_all0.record(stream)
    for i in range(args.steps):
        if i % FRAME_SKIP == 0:
            ctrl_buf.uniform_(0.0, 1.0, generator=gen)
        ev[i][0].record(stream)
        mjx.step(model, d, args.nsub)
        ev[i][1].record(stream)
    d.join()              # pipelined outputs: the last step's wide-output kernel ends inside the timed region
    t_all1.record(stream)
    torch.cuda.synchronize()
    if world > 1:
        dist.barrier()
    clocks = sampler.stop()
    launches = lib.lib().myo_launch_count() - launches0
    total_ms = t_all0.elapsed_time(t_all1)
    kernel_ms = sum(a.elapsed_time(b) for a, b in ev) / args.steps
    if args.pipelined:    # the output kernel overlaps the next step: per-step time = timed region / steps
        kernel_ms = total_ms / args.steps
    tt = torch.tensor([total_ms], device=dev)
    if world > 1:
        dist.all_reduce(tt, op=dist.ReduceOp.MAX)
    total_ms = float(tt.item())
    env_steps = nw * world * args.steps * args.nsub
    value = env_steps / (total_ms * 1e-3)

    # ---- end to end through the C ABI with host buffers (pinned), every step: H2D of the step's input (ctrl; the state is
    # carried on the device from the previous step, as in the reference's env.step(a) loop), step, D2H of the new state
    # host buffers in the ABI layout [k][nworld]
    hq = torch.empty((m.nq, nw), dtype=torch.float32).pin_memory(); hv = torch.zeros((m.nv, nw), dtype=torch.float32).pin_memory()
    ha = torch.zeros((m.na, nw), dtype=torch.float32).pin_memory(); hc = torch.rand((m.nu, nw), dtype=torch.float32).pin_memory()
    hq.copy_(d.storage("qpos").cpu()); hv.copy_(d.storage("qvel").cpu()); ha.copy_(d.storage("act").cpu())
    import ctypes
    L = lib.lib()
    vp = ctypes.c_void_p
    e2e_steps = max(10, args.steps // 4)

    def host_step():
        lib.check(L.myo_step_host(d._h, args.nsub, None, None, None, vp(hc.data_ptr()),
                                  vp(hq.data_ptr()), vp(hv.data_ptr()), vp(ha.data_ptr()), vp(stream.cuda_stream)))
    def e2e_run():
        for _ in range(3):
            host_step()
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(stream)
        for _ in range(e2e_steps):
            host_step()          # H2D ctrl (pinned) -> step -> D2H qpos / qvel / act, synchronises the stream
        d.join()                 # pipelined outputs: the last wide-output kernel ends inside the timed region
        e1.record(stream)
        torch.cuda.synchronize()
        te = torch.tensor([e0.elapsed_time(e1)], device=dev)
        if world > 1:
            dist.all_reduce(te, op=dist.ReduceOp.MAX)
        return nw * world * e2e_steps * args.nsub / (float(te.item()) * 1e-3)

    # A host-buffer caller receives the state; the wide mjData fields stay on the device.  First with every field complete
    # in stream order when the call returns (as `value` above), then with the opt-in pipelined outputs: the wide-output
    # kernel of a step runs on the library's side stream under the next call's upload and step kernel.
    state0 = {k: d.storage(k).clone() for k in ("qpos", "qvel", "act", "qacc_warmstart", "time")}   # both runs start from the same state
    e2e_ordered = e2e_run()
    e2e_value = e2e_ordered
    if args.fields == "all" and not args.pipelined:
        for k, v in state0.items():
            d.storage(k).copy_(v)
        d.set_pipelined(True)
        e2e_value = e2e_run()
        d.set_pipelined(False)
    h2d = nw * m.nu * 4
    d2h = nw * (m.nq + m.nv + m.na) * 4

    # ---- optional episode statistics all-reduce: the only collective, off the step path
    stats = torch.stack([d.status.ne(0).sum().float(), d.qvel.abs().max().float()])
    if world > 1:
        dist.all_reduce(stats[:1])
    if rank != 0:
        if world > 1:
            dist.destroy_process_group()
        return

    # ---- roofline of the fused step kernel (HBM bound by construction of the metric: bytes of every bound field)
    bytes_per_world = 0
    for name in d.fields():
        t = getattr(d, name)
        bytes_per_world += t[0].numel() * 4 * (2 if name in ("qpos", "qvel", "act", "time", "qacc_warmstart", "status") else 1)
    peaks_path = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(peaks_path):
        peak, peak_src = json.load(open(peaks_path))["hbm_gbs"], "measured"
    else:
        peak, peak_src = 6650.0, "fallback"
    achieved = bytes_per_world * nw * args.nsub / (kernel_ms * 1e-3) / 1e9 if args.nsub == 1 else bytes_per_world * nw / (kernel_ms * 1e-3) / 1e9
    traffic = None
    tp = os.path.join(ROOT, "profiles", "traffic.json")
    if os.path.exists(tp):
        traffic = json.load(open(tp)).get(f"{args.model}_{args.fields}_{nw}")
    line = {"metric": "env-steps/s", "value": value, "unit": "env-steps/s", "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
            "ms_per_step": total_ms / args.steps, "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f32",
            "data": "synthetic",
            "config": {"workload": WORKLOAD.format(model=args.model, nworld=nw), "nworld_per_gpu": nw, "fields": args.fields,
                       "nsub_per_launch": args.nsub, "frame_skip_ctrl_hold": FRAME_SKIP, "nconmax": d.nconmax, "njmax": d.njmax,
                       "scratch_bytes_per_world": d.smem_bytes, "lanes_per_world": d.team, "step_ctas_per_sm": d.ctas_per_sm, "pipelined_outputs": bool(args.pipelined), "bytes_per_world_step": bytes_per_world,
                       "l2": "working set per step (%.0f MB) exceeds the 126 MB L2" % (bytes_per_world * nw / 1e6),
                       "partition": "worlds by index, no collective on the step path",
                       "worlds_with_status_flags": int(stats[0].item())},
            "clocks": clocks, "gpu_launches": int(launches),
            "e2e": {"value": e2e_value, "unit": "env-steps/s", "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": d2h,
                    "api": "myo_step_host (C ABI, pinned host buffers: ctrl up, qpos / qvel / act down, every step)" +
                           ("; wide-output kernel pipelined under the next call (myo_batch_set_pipelined)" if e2e_value != e2e_ordered or args.pipelined else ""),
                    "stream_ordered_value": e2e_ordered, "steps": e2e_steps},
            "roofline": {"bound": "hbm", "achieved": achieved, "peak": peak, "unit": "GB/s", "frac": achieved / peak, "traffic": traffic,
                         "peak_source": peak_src, "kernel": ("stepKernel (team.cu, %d lanes per world) + outKernel (out.cu): one myo_step" if args.fields == "all" else "stepKernel (team.cu, %d lanes per world)") % d.team,
                         "kernel_ms": kernel_ms,
                         "algorithmic_bytes_per_launch": bytes_per_world * nw}}
    if not args.no_cpu:
        line["cpu_baseline"] = cpu_baseline(m, os.cpu_count() or 1)
    print(json.dumps(line), flush=True)
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
