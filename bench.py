#!/usr/bin/env python
"""Headline benchmark: env-steps/s of the batched MyoHand physics step (BASELINE.json metric).

  python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference]
  python -m torch.distributed.run --nnodes=1 --nproc-per-node N ... bench.py --gpus N ...

One "step" = one mj_step of every world of this rank (one myo_step call = the phase pipeline of csrc/tile.cu), on the
MyoHand model of myosuite/mjx/play.py:8 with 8192 worlds per GPU, random ctrl ~ U[0,1) redrawn every 10 steps
(SURVEY.md 8d).  Worlds are partitioned by index; start states and controls are functions of the GLOBAL world index
(myosuite_mjx_b200/partition.py), so a world's trajectory does not depend on the GPU count.  Prints ONE JSON line (rank 0).

`--impl reference` times the CPU arm on the host cores instead: real MuJoCo `mj_step` when `import mujoco` succeeds and
the MJCF files are reachable (oracle/mujoco_probe.py), else the in-repo CPU restatement (oracle/): the reference's own
CPU path (MuJoCo C through dm_control) is third-party code that is neither vendored nor installed here.
PARITY of the physics step is UNPINNED (no MuJoCo output available to pin the oracle): the JSON line says so.
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
STEADY_AFTER = 300      # rollout steps after which the hand has clenched (13 contacts / 65 rows per world): "steady state"
STEADY_STEPS = 100
PARITY = "unpinned: checked against the in-repo CPU restatement of MuJoCo 3.2.8 semantics (oracle/), no MuJoCo output available"


def load_model(name):
    from myosuite_mjx_b200 import mjcf
    return mjcf.Model.load(os.path.join(ROOT, "myosuite_mjx_b200", "models", name + ".npz"))


def initial_qpos(m, nworld, seed, w0=0):
    """qpos0 + U(-0.1, 0.1) * range, clipped to the joint range, per GLOBAL world index."""
    from myosuite_mjx_b200 import partition
    return partition.initial_qpos(m, range(w0, w0 + nworld), seed)


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


class CpuArm:
    """The CPU implementation of the path on the host cores: a CONTINUING random-ctrl rollout of `nworld` worlds (ctrl held
    FRAME_SKIP steps), one `advance(n)` = n mj_steps of every world on all threads.  MuJoCo itself when importable, else the
    oracle port."""

    def __init__(self, m, model_name, cores):
        sys.path.insert(0, os.path.join(ROOT, "oracle"))
        import mujoco_probe
        self.cores, self.m = cores, m
        self.nworld = 4 * cores
        q0 = initial_qpos(m, self.nworld, SEED)
        self.mj = mujoco_probe.rollout_backend(model_name, self.nworld, cores, q0, SEED)
        self.kind = "mujoco" if self.mj is not None else "port"
        if self.mj is None:
            import oracle
            self.pool = oracle.Pool(m, self.nworld, q0, SEED)

    def advance(self, nsteps):
        if self.mj is not None:
            self.mj.advance(nsteps, FRAME_SKIP)
        else:
            self.pool.advance(nsteps, FRAME_SKIP, self.cores)
        return self.nworld * nsteps

    def describe(self, nsteps_total):
        what = "MuJoCo mj_step (python bindings, one MjData per world, thread pool)" if self.kind == "mujoco" else \
            "float64 CPU restatement of mj_step (oracle/), NOT MuJoCo"
        return f"{self.nworld} worlds x {nsteps_total} mj_steps of one continuing rollout, random ctrl held {FRAME_SKIP} steps, {self.cores} threads, {what}"


def cpu_baseline(m, model_name, cores, target_seconds=12.0):
    """CPU arm over a bounded sample of the same workload (about `target_seconds` of host work)."""
    arm = CpuArm(m, model_name, cores)
    t = time.time()
    arm.advance(50)                                # calibrate (and leave the start pose)
    rate = arm.nworld * 50 / max(time.time() - t, 1e-6)
    nsteps = int(max(50, min(2000, target_seconds * rate / arm.nworld)))
    t = time.time()
    n = arm.advance(nsteps)
    dt = time.time() - t
    return {"value": n / dt, "unit": "env-steps/s", "cores": cores, "kind": arm.kind, "sample": arm.describe(nsteps)}


def run_reference(args, rank, world):
    """CPU arm on every host thread (rank 0 only): each bench "step" is 20 mj_steps of the continuing rollout."""
    if rank != 0:
        return
    m = load_model(args.model)
    cores = os.cpu_count() or 1
    arm = CpuArm(m, args.model, cores)
    per_step = 20
    for _ in range(args.warmup):
        arm.advance(per_step)
    t = time.time()
    n = 0
    for _ in range(args.steps):
        n += arm.advance(per_step)
    dt = time.time() - t
    val = n / dt
    sample = arm.describe(per_step * args.steps) + f" ({per_step} mj_steps per bench step)"
    line = {"impl": "reference", "metric": "env-steps/s", "value": val, "unit": "env-steps/s", "n_gpus": args.gpus, "steps": args.steps,
            "warmup": args.warmup, "ms_per_step": 1e3 * dt / args.steps, "higher_is_better": True, "scaling": "weak",
            "vs_baseline": None, "dtype": "f64", "data": "synthetic", "parity": PARITY,
            "config": {"workload": f"CPU arm of the MyoHand ({args.model}.xml) random-ctrl rollout: {sample}; the GPU arm runs {args.nworld} worlds/GPU"},
            "cpu_baseline": {"value": val, "unit": "env-steps/s", "cores": cores, "kind": arm.kind, "sample": sample},
            "e2e": {"value": val, "unit": "env-steps/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}}
    print(json.dumps(line), flush=True)


WORKLOAD = "MyoHand ({model}.xml, play.py:8) random-ctrl rollout, {nworld} worlds/GPU, dt 0.002, Newton, full mjData written every step"
KERNELS = "one myo_step = kin, then collide > collideMpr > collideCompact > makecon | tendon > velact | inertia | tenout, conout (output streams), then newton (+ mj_Euler tail) | posout: 12 kernels + 2 empty ordering launches (csrc/tile.cu, lane = world)"


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
    ap.add_argument("--no-steady", action="store_true", help="skip the steady_state leg")
    ap.add_argument("--count-aware", action="store_true", help="opt-in myo_batch_set_count_aware: contact_* / efc_* written up to ncon / nefc only (NOT the headline configuration)")
    args = ap.parse_args()
    rank = int(os.environ.get("RANK", "0")); world = int(os.environ.get("WORLD_SIZE", "1")); local = int(os.environ.get("LOCAL_RANK", "0"))
    if args.impl == "reference":
        run_reference(args, rank, world)
        return

    import torch
    import torch.distributed as dist
    from myosuite_mjx_b200 import lib, mjx, partition
    if not torch.cuda.is_available():
        raise SystemExit("bench.py needs a CUDA device: the product has no CPU fallback")
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)
    m = load_model(args.model)
    model = mjx.put_model(m, device=local)
    nw = args.nworld
    # worlds are partitioned by index: rank r owns [r*nw, (r+1)*nw); start states and controls are functions of the global index
    w0 = rank * nw
    d = mjx.put_data(model, dict(qpos=initial_qpos(m, nw, SEED, w0)), nworld=nw, fields=args.fields)
    if args.count_aware:
        d.set_count_aware(True)
    ctrl_buf = d.storage("ctrl")
    stream = torch.cuda.current_stream()
    step_no = [0]

    def one_step():
        if step_no[0] % FRAME_SKIP == 0:
            ctrl_buf.copy_(partition.ctrl_hash_torch(m.nu, w0, nw, SEED, step_no[0] // FRAME_SKIP, dev))
        mjx.step(model, d, args.nsub)
        step_no[0] += 1

    def timed(nsteps):
        """nsteps steps bracketed by barrier + synchronize; device time of the region (max over ranks) and mean per-step kernel time"""
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()
        ev = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(nsteps)]
        t0, t1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        t0.record(stream)
        for i in range(nsteps):
            if step_no[0] % FRAME_SKIP == 0:
                ctrl_buf.copy_(partition.ctrl_hash_torch(m.nu, w0, nw, SEED, step_no[0] // FRAME_SKIP, dev))
            ev[i][0].record(stream)
            mjx.step(model, d, args.nsub)
            ev[i][1].record(stream)
            step_no[0] += 1
        t1.record(stream)
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
        tt = torch.tensor([t0.elapsed_time(t1)], device=dev)
        if world > 1:
            dist.all_reduce(tt, op=dist.ReduceOp.MAX)
        return float(tt.item()), sum(a.elapsed_time(b) for a, b in ev) / nsteps

    for _ in range(args.warmup):
        one_step()
    sampler = ClockSampler(local); sampler.start()
    launches0 = lib.lib().myo_launch_count()
    total_ms, kernel_ms = timed(args.steps)
    clocks = sampler.stop()
    launches = lib.lib().myo_launch_count() - launches0
    value = nw * world * args.steps * args.nsub / (total_ms * 1e-3)

    # ---- end to end through the C ABI with pinned host buffers, every step: H2D of the step's input (ctrl; the state is
    # carried on the device from the previous step, as in the reference's env.step(a) loop), step, D2H of the new state.
    # Host buffers in the ABI layout [k][nworld].  The full mjData stays on the device (a host caller receives the state).
    # Runs right after the device-timed leg, on the same rollout (same ctrl sequence), so `e2e` and `value` see the same phase.
    hq = torch.empty((m.nq, nw), dtype=torch.float32).pin_memory(); hv = torch.zeros((m.nv, nw), dtype=torch.float32).pin_memory()
    ha = torch.zeros((m.na, nw), dtype=torch.float32).pin_memory()
    import ctypes
    L = lib.lib()
    vp = ctypes.c_void_p
    e2e_steps = max(10, min(250, args.steps // 4))

    # the policy's outputs for the leg, one pinned [nu][nworld] buffer per control frame (made before the timed region)
    frame0 = step_no[0] // FRAME_SKIP
    nframes = (step_no[0] + e2e_steps + 3) // FRAME_SKIP - frame0 + 1
    hcs = [torch.from_numpy(np.ascontiguousarray(partition.ctrl_hash_np(m.nu, range(w0, w0 + nw), SEED, frame0 + f).T.astype(np.float32))).pin_memory()
           for f in range(nframes)]

    def host_step():
        hc = hcs[step_no[0] // FRAME_SKIP - frame0]
        lib.check(L.myo_step_host(d._h, args.nsub, None, None, None, vp(hc.data_ptr()),
                                  vp(hq.data_ptr()), vp(hv.data_ptr()), vp(ha.data_ptr()), vp(stream.cuda_stream)))
        step_no[0] += 1
    for _ in range(3):
        host_step()
    torch.cuda.synchronize()
    if world > 1:
        dist.barrier()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record(stream)
    for _ in range(e2e_steps):
        host_step()          # H2D ctrl (pinned) -> step -> D2H qpos / qvel / act, synchronises the stream
    e1.record(stream)
    torch.cuda.synchronize()
    te = torch.tensor([e0.elapsed_time(e1)], device=dev)
    if world > 1:
        dist.all_reduce(te, op=dist.ReduceOp.MAX)
    e2e_value = nw * world * e2e_steps * args.nsub / (float(te.item()) * 1e-3)
    h2d = nw * m.nu * 4
    d2h = nw * (m.nq + m.nv + m.na) * 4

    # ---- steady state: the same rollout once the hand has clenched (a short driver run times the contact-poor first steps)
    steady = None
    if not args.no_steady:
        while step_no[0] < STEADY_AFTER:
            one_step()
        after = step_no[0]
        s_ms, s_kernel_ms = timed(STEADY_STEPS)
        fields = d.fields()
        steady = {"value": nw * world * STEADY_STEPS * args.nsub / (s_ms * 1e-3), "unit": "env-steps/s", "ms_per_step": s_ms / STEADY_STEPS,
                  "kernel_ms": s_kernel_ms, "after_steps": after, "steps": STEADY_STEPS,
                  "ncon_mean": float(d.ncon.float().mean()) if "ncon" in fields else None,
                  "nefc_mean": float(d.nefc.float().mean()) if "nefc" in fields else None}

    # ---- optional episode statistics all-reduce: the only collective, off the step path
    stats = partition.all_reduce_stats(partition.episode_stats(d.status, d.qvel))
    if rank != 0:
        if world > 1:
            dist.destroy_process_group()
        return

    # ---- roofline of one myo_step (HBM bound by construction of the metric: bytes of every bound field, written once per step,
    # the state read and written).  `live`: the same count with contact_* / efc_* taken up to the mean ncon / nefc instead of
    # their zero-filled capacity (MJX pads the same way; mjData does not).
    bytes_per_world, live_bytes = 0, 0
    per_con = {"contact_geom": 2, "contact_dim": 1, "contact_efc_address": 1, "contact_dist": 1, "contact_pos": 3, "contact_frame": 9,
               "contact_includemargin": 1, "contact_friction": 5, "contact_solref": 2, "contact_solimp": 5}
    per_row = {"efc_j": m.nv, "efc_pos": 1, "efc_margin": 1, "efc_d": 1, "efc_aref": 1, "efc_force": 1, "efc_type": 1}
    nc = steady["ncon_mean"] if steady and steady["ncon_mean"] is not None else 0.0
    ne = steady["nefc_mean"] if steady and steady["nefc_mean"] is not None else 0.0
    for name in d.fields():
        n = getattr(d, name)[0].numel() * 4 * (2 if name in ("qpos", "qvel", "act", "time", "qacc_warmstart", "status") else 1)
        bytes_per_world += n
        live_bytes += per_con[name] * 4 * nc if name in per_con else (per_row[name] * 4 * ne if name in per_row else n)
    peaks_path = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(peaks_path):
        peak, peak_src = json.load(open(peaks_path))["hbm_gbs"], "measured"
    else:
        peak, peak_src = 6650.0, "fallback"
    achieved = bytes_per_world * nw / (kernel_ms / args.nsub * 1e-3) / 1e9
    traffic = None
    tp = os.path.join(ROOT, "profiles", "traffic.json")
    if os.path.exists(tp):
        traffic = json.load(open(tp)).get(f"{args.model}_{args.fields}_{nw}")
    line = {"metric": "env-steps/s", "value": value, "unit": "env-steps/s", "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
            "ms_per_step": total_ms / args.steps, "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f32",
            "data": "synthetic", "parity": PARITY,
            "config": {"workload": WORKLOAD.format(model=args.model, nworld=nw) + (" (count-aware: contact / constraint records up to ncon / nefc only)" if args.count_aware else ""),
                       "nworld_per_gpu": nw, "fields": args.fields, "count_aware": bool(args.count_aware),
                       "nsub_per_launch": args.nsub, "frame_skip_ctrl_hold": FRAME_SKIP, "nconmax": d.nconmax, "njmax": d.njmax,
                       "scratch_bytes_per_world": d.smem_bytes, "engine": "phase pipeline, lane = world (32 worlds per tile CTA)",
                       "bytes_per_world_step": bytes_per_world, "live_bytes_per_world_step": int(live_bytes),
                       "l2": "working set per step (%.0f MB) exceeds the 126 MB L2" % (bytes_per_world * nw / 1e6),
                       "partition": "worlds by index, no collective on the step path; seeds by global world index",
                       "worlds_with_status_flags": int(stats[1].item())},
            "clocks": clocks, "gpu_launches": int(launches),
            "steady_state": steady,
            "e2e": {"value": e2e_value, "unit": "env-steps/s", "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": d2h,
                    "api": "myo_step_host (C ABI, pinned host buffers: ctrl up, qpos / qvel / act down, every step; stream-ordered, the default API behaviour)",
                    "steps": e2e_steps},
            "roofline": {"bound": "hbm", "achieved": achieved, "peak": peak, "unit": "GB/s", "frac": achieved / peak, "traffic": traffic,
                         "peak_source": peak_src, "kernel": KERNELS, "kernel_ms": kernel_ms,
                         "algorithmic_bytes_per_launch": bytes_per_world * nw,
                         "steady_state_frac": (bytes_per_world * nw / (steady["kernel_ms"] / args.nsub * 1e-3) / 1e9 / peak) if steady else None,
                         "live_frac": (live_bytes * nw / (steady["kernel_ms"] / args.nsub * 1e-3) / 1e9 / peak) if steady else None}}
    if not args.no_cpu:
        line["cpu_baseline"] = cpu_baseline(m, args.model, os.cpu_count() or 1)
    print(json.dumps(line), flush=True)
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
