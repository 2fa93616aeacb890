#!/usr/bin/env python
"""bench.py - predicted node-timesteps/s of the Emulator rollout (BASELINE.json metric).

  python bench.py --gpus 1 --steps K --warmup W            # this repo's CUDA path
  python bench.py --impl reference --gpus 1 --steps K ...  # the reference math on host cores
  torchrun ... bench.py --gpus N ...                       # one rank per GPU, weak scaling

A step = one Emulator.predict over one batch of synthetic scenarios.  At N=1 the workload is
BASELINE.json configs[1] (C2: shunqing GAT+Conv1D edge_fusion/resnet/flood, 60->60, B=4096).
`value`   : inputs already resident in HBM, device-side timing (CUDA events), max over ranks.
`e2e`     : the same through Emulator.predict with pinned HOST buffers (H2D + D2H inside).
`e2e`     : the same through Emulator.predict with pinned HOST buffers (H2D + D2H inside); `e2e.numpy_api` is the
            caller's real entry, Emulator.predict(states, b, a, edge_state) on ordinary (pageable) NumPy arrays.
`roofline`: the dominant kernel family timed live with CUDA events on the library's stream.  Small-graph models
            (graph_base 0, C1-C4) are dense contractions -> bound "tensor" (SURVEY 8d): achieved = the FLOPs the fused
            spatial kernel executes per launch / its average launch time, peak = bf16_tflops_sustained.  The
            layer-by-layer HBM view of the same kernel is kept as `design_hbm_frac`; `io_bytes_per_step` is 8(d)'s
            algorithmic I/O (inputs + outputs), `traffic` the ncu dram bytes per launch.  graph_base 1 (C5): the CSR
            gather, bound "hbm".
`cpu_baseline`: the NumPy oracle (a port - the reference needs TF/spektral, absent here) on a
                bounded sample of the same workload, rank 0, N=1 only; `cpu_baseline_c1` is C1 exactly as
                BASELINE.json words it (astlingen GCN+GRU 10->10, B=64), median of 5.
"""
import argparse
import json
import os
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))

METRIC = "predicted node-timesteps/sec (shunqing GAT rollout)"
UNIT = "node-timesteps/s"


def flops(em, B):
    """Algorithmic FLOPs (2*MAC, CSR aggregation) of one predict; SURVEY.md 8(d) formula.
    Returns (total, gemm_part) - gemm_part is what the rows-GEMM kernel family executes."""
    E, H, S, P, k = em.embed_size, em.hidden_dim, em.n_sp_layer, em.n_tp_layer, em.kernel_size
    N, M, Ti, To, F = em.n_node, em.n_edge, em.seq_in, em.seq_out, em.if_flood
    act = 1 if em.act else 0
    nnz_n, nnz_m = em.filter_csr.nnz, (em.edge_filter_csr.nnz if em.edge_filter_csr is not None else 0)
    gat = em.conv_kind == "gat"
    tot = 2 * (B * Ti * (N * em.n_in + M * em.e_in) * E + B * To * (N * em.b_in + M * act) * E // 2)
    gemm = 0
    if em.conv_kind == "nn":       # NN-cat: one flattened row per (b,t) (emulator.py:231-233), everything is a row GEMM
        for blk, T in ((0, Ti), (1, To)):
            c = 2 * E if blk == 0 else 2 * H + E // 2 + (E // 2) * act
            for l in range(S):
                tot += 2 * B * T * c * 2 * E
                c = 2 * E
            for l in range(P):
                c = E if l == 0 else H
                tot += 2 * 2 * B * T * {"Conv1D": k * c * H, "GRU": 3 * H * (c + H), "LSTM": 4 * H * (c + H)}[em.recurrent]
        tot += 2 * B * To * (2 * H * E + (E * E // 2 + (F - 1) * (E // 2) ** 2 + (E // 2) * N if F else 0)
                             + E * em.n_out * N + E * em.e_out * M)
        return float(tot), float(tot)
    for blk, T in ((0, Ti), (1, To)):
        cx, ce = (E, E) if blk == 0 else (H + E // 2, H + (E // 2) * act)
        for l in range(S):
            if em.graph_base:      # one conv over the joint node+edge graph (emulator.py:220-223)
                g = (N * cx + M * ce) * E
                o = (2 * (N + M) * E if gat else 0) + nnz_n * E
            else:
                g = M * ce * (E // 2) + N * cx * (E // 2) + N * (cx + E // 2) * E + M * (ce + E // 2) * E
                o = 2 * N * M * (E // 2) + (2 * (N + M) * E if gat else 0) + (nnz_n + nnz_m) * E
            gemm += 2 * B * T * g
            tot += 2 * B * T * (g + o)
            cx = ce = E
        for l in range(P):
            c = E if l == 0 else H
            per = {"Conv1D": k * c * H, "GRU": 3 * H * (c + H), "LSTM": 4 * H * (c + H)}[em.recurrent]
            gpart = {"Conv1D": k * c * H, "GRU": 3 * H * c, "LSTM": 4 * H * c}[em.recurrent]
            tot += 2 * B * (N + M) * T * per
            gemm += 2 * B * (N + M) * T * gpart
    heads_g = N * H * E + M * H * E + (N * (E * E // 2 + (F - 1) * (E // 2) ** 2) if F else 0)
    heads_o = N * E * em.n_out + M * E * em.e_out + (N * (E // 2) if F else 0)
    tot += 2 * B * To * (heads_g + heads_o)
    gemm += 2 * B * To * heads_g
    return float(tot), float(gemm)


def dominant_kernel(em, B, precision):
    """Which kernel family dominates a step and its ALGORITHMIC bytes per step.
    graph_base=0, tensor-core path: sp_conv_kernel (fused NodeEdge+conv+aggregation), HBM-bound by its
    arithmetic intensity (~100 FLOP/B < 212 FLOP/B machine balance): per sample and side it must read
    self (R x Kself) + other_dense (Ro x E/2) and write out (R x E), bf16.
    graph_base=1: agg_large_kernel (CSR gather): nnz*E gathered + R*E written per sample and layer.
    fp32 path: the rows-GEMM family (tensor roofline kept for reference)."""
    E, H, S = em.embed_size, em.hidden_dim, em.n_sp_layer
    N, M, Ti, To = em.n_node, em.n_edge, em.seq_in, em.seq_out
    es = 2 if precision == "bf16" else 4
    if em.graph_base:
        R, nnz = N + M, em.filter_csr.nnz
        per_sample = (nnz * E + R * E) * es
        return dict(family=2, kernel="agg_large16_kernel (CSR neighbour gather, joint node+edge graph; agg_large_kernel on the fp32 path)" if precision == "bf16" else "agg_large_kernel (CSR neighbour gather, joint node+edge graph)", bound="hbm",
                    bytes=float(B * (Ti + To) * S * per_sample), unit_note="(nnz*E gathered + (N+M)*E written) * %d B per (b,t) sample and layer" % es)
    if precision != "bf16" or em.conv_kind == "nn":
        return dict(family=0, kernel="rows_gemm_f32 (all Dense/conv-linear/Conv1D launches of a step)", bound="fp32", bytes=None, unit_note="")
    act = 1 if em.act else 0
    gat = em.conv_kind == "gat"
    nnz_n, nnz_m = em.filter_csr.nnz, em.edge_filter_csr.nnz
    tot, fl = 0, 0
    for blk, T in ((0, Ti), (1, To)):
        cx, ce = (E, E) if blk == 0 else (H + E // 2, H + (E // 2) * act)
        for l in range(S):
            tot += B * T * ((N * cx + M * (E // 2) + N * E) + (M * ce + N * (E // 2) + M * E)) * 2
            # what the fused kernel executes (2*MAC): NodeEdge both sides, conv linear part, GAT scores, CSR aggregation
            fl += 2 * B * T * (2 * N * M * (E // 2) + N * (cx + E // 2) * E + M * (ce + E // 2) * E
                               + (2 * (N + M) * E if gat else 0) + (nnz_n + nnz_m) * E)
            cx = ce = E
    io = 4 * B * (Ti * (N * em.n_in + M * em.e_in) + To * (N * em.b_in + (em.n_act if em.act else 0))
                  + To * (N * em.y_channels + M * em.ey_channels))
    return dict(family=1, kernel="sp_agg_kernel / sp_conv_kernel (fused NodeEdge + conv + GAT/GCN aggregation, both stream sides)",
                bound="tensor", bytes=float(tot), flops=float(fl), io_bytes=float(io),
                unit_note="FLOPs per sample and side: 2*(R*Ro*E/2 + R*(Kself+E/2)*E [+ 2*R*E GAT scores] + nnz*E); "
                          "design_hbm bytes: (R*Kself + Ro*E/2 read + R*E written) * 2 B")


class ClockSampler:
    """nvidia-smi clocks + throttle reasons sampled during the timed region (B200_PROFILING.md)."""
    Q = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, gpu):
        self.gpu, self.rows, self.proc = gpu, [], None

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", "-i", str(self.gpu), "--query-gpu=" + self.Q,
                                          "--format=csv,noheader,nounits", "-lms", "100"],
                                         stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.th = threading.Thread(target=self._read, daemon=True)
            self.th.start()
        except Exception:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append([c.strip() for c in line.split(",")])

    def stop(self):
        if not self.proc:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        self.proc.terminate()
        try:
            self.proc.wait(timeout=5)
        except Exception:
            self.proc.kill()
        sm, mx, reasons = [], [], set()
        for r in self.rows:
            try:
                sm.append(float(r[1])); mx.append(float(r[2]))
                for name, v in zip(("hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"), r[5:9]):
                    if v.lower().startswith("active"):
                        reasons.add(name)
            except Exception:
                pass
        return {"sm_mhz": float(np.median(sm)) if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "reasons": sorted(reasons), "samples": len(sm)}


def peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        d = json.load(open(p))
        return d.get("bf16_tflops_sustained", 1399.6), d.get("hbm_gbs", 6554.6), "measured"
    return 1400.0, 6650.0, "fallback"


def traffic_per_launch(dom, workload, B, launches_per_step):
    """dram__bytes_read + dram__bytes_write per launch of the dominant kernel from the committed ncu capture
    (profiles/r02_traffic.json: bytes per scenario and step, measured at one batch size; the kernels stream
    scenarios independently, so the figure scales linearly with B)."""
    p = os.path.join(ROOT, "profiles", "r02_traffic.json")
    if not os.path.exists(p):
        return None
    d = json.load(open(p)).get(workload, {}).get(dom["kernel"].split(" ")[0])
    if not d:
        return None
    return d["dram_bytes_per_scenario_step"] * B / launches_per_step


def cpu_baseline(workload, seconds=20.0, fixed_B=None, runs=3):
    """NumPy oracle (port) fp32 on a bounded sample of the workload (or exactly fixed_B scenarios); all BLAS threads."""
    import cases   # noqa: F401  (tests/cases.py helpers are not needed; keep path set up)
    from gnn_uds_b200 import configs, synthetic, weights as Wt
    from oracle.emulator_oracle import OracleEmulator
    args = configs.make_args(workload)
    o = OracleEmulator(args.conv, args.resnet, args.recurrent, args, dtype=np.float32)
    shim = cases.SpecShim(o)
    o.set_weights(Wt.init_weights(Wt.layer_specs(shim), args.seed, node_edge_bias_std=0.05, bias_std=0.05))
    rng = np.random.default_rng(args.seed)
    n = synthetic.make_norms(o.n_node, o.n_edge, o.if_flood, rng, tide=o.tide)
    o.set_norm(n["norm_x"], n["norm_b"], n["norm_y"], n["norm_r"], n["norm_e"])
    shim.norm_x, shim.norm_b, shim.norm_e = n["norm_x"], n["norm_b"], n["norm_e"]
    Bs = 1
    inp = synthetic.make_inputs(shim, Bs, rng)
    t0 = time.perf_counter(); o.predict(*inp); t1 = time.perf_counter() - t0      # warm-up + calibration
    Bs = int(fixed_B) if fixed_B else int(max(1, min(64, seconds / 3.0 / max(t1, 1e-3))))
    inp = synthetic.make_inputs(shim, Bs, rng)
    ts = []
    for _ in range(runs):
        t0 = time.perf_counter(); o.predict(*inp); ts.append(time.perf_counter() - t0)
    t = float(np.median(ts))
    return {"value": Bs * o.seq_out * o.n_node / t, "unit": UNIT, "cores": os.cpu_count(), "kind": "port",
            "sample": "%s model, B=%d scenarios x %d steps, NumPy fp32 dense formulation, median of %d" % (workload, Bs, o.seq_out, runs),
            "ms_per_sample_call": t * 1e3}


def reference_config(a):
    """The b200 arm's `config` for the same workload (same wording), plus what the CPU arm actually ran per step."""
    from gnn_uds_b200 import configs
    args = configs.make_args(a.workload)
    conv = "NN" if str(args.conv) in ("False", "None") else ("GAT" if "GAT" in str(args.conv) else "GCN")
    N, M = len(args.nodes), len(args.links)
    B = a.batch or args.B
    return {"workload": "%s: %s %s+%s edge_fusion=%s resnet=%s if_flood=%d act=%s, N=%d M=%d, seq %d->%d, B=%d per GPU" % (
                a.workload, configs.WORKLOADS[a.workload]["net"], conv, args.recurrent, bool(getattr(args, "edge_fusion", False)),
                bool(args.resnet), int(getattr(args, "if_flood", 0)), bool(getattr(args, "act", False)), N, M, args.seq_in, args.seq_out, B),
            "precision": "fp32", "sharding": "rank 0 only: host cores",
            "sample": "each step = a bounded sample of the workload (B of a few scenarios, see cpu_baseline.sample); the rate is per node-timestep, so it compares with the full batch"}


def run_reference(a):
    """--impl reference: the reference's math on host cores (oracle port; TF/spektral are absent)."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    ts = []
    res = None
    for i in range(a.warmup + a.steps):
        res = cpu_baseline(a.workload, seconds=6.0)
        if i >= a.warmup:
            ts.append(res["value"])
    v = float(np.mean(ts))
    res["value"] = v
    line = {"impl": "reference", "metric": METRIC, "value": v, "unit": UNIT, "n_gpus": a.gpus, "steps": a.steps,
            "warmup": a.warmup, "ms_per_step": res["ms_per_sample_call"], "higher_is_better": True, "scaling": "weak",
            "vs_baseline": None, "dtype": "f32", "data": "synthetic", "config": reference_config(a),
            "cpu_baseline": res, "e2e": {"value": v, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}}
    emit(line)


def run_mpc(a):
    """--workload C3mpc: BASELINE.json configs[2] as north_star words it - an 8192-candidate population (TOTAL, strong
    scaling) x 60-step horizon = 12 autoregressive chunks of the astlingen act+edge_fusion+flood 5->5 model, candidates
    sharded over the ranks, evaluated by MpcEvaluator.pred_emu_device: only the decision matrix goes in, the rollout,
    the feedback rule and the objective stay on the device, and every step ends with the ONE collective of the path,
    the NCCL all-gather of the per-candidate objectives."""
    import torch
    import torch.distributed as dist
    from gnn_uds_b200 import configs, synthetic, mpc, sharding
    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)
    precision = "bf16" if a.precision == "auto" else a.precision
    em, args, rng = configs.make_emulator("C3b", device=local, precision=precision)
    net, _ = configs.load_net("astlingen")
    pop = a.batch or 8192
    chunks = configs.WORKLOADS["C3b"]["chunks"]
    T = em.seq_out * chunks
    ev = mpc.MpcEvaluator(em, net.nodes, interval=1, eval_horizon=T, control_horizon=T, setting_duration=5,
                          gamma=0.99 ** np.arange(T))
    rng = np.random.default_rng(args.seed)          # same state and population on every rank
    states, b, _, ex = synthetic.make_inputs(em, 1, rng, horizon=T)
    ev.load_state(states[0], b[0], ex[0])
    x = mpc.latin_hypercube(pop, np.full(ev.n_var, 0.05), np.ones(ev.n_var), rng)

    def barrier():
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    for _ in range(max(a.warmup, 3)):
        F = ev.pred_emu_device(x)
    barrier()
    sampler = ClockSampler(local)
    if rank == 0:
        sampler.start()
    em.set_profiling(1)
    l0 = em.launch_count()
    dev_ms = 0.0
    t0 = time.perf_counter()
    for _ in range(a.steps):
        F = ev.pred_emu_device(x)
        dev_ms += em.last_timing()[0]
    barrier()
    wall = time.perf_counter() - t0
    launches = em.launch_count() - l0
    clocks = sampler.stop() if rank == 0 else None
    # the collective alone (CUDA events on torch's stream): per-candidate objectives of every shard
    gather_us = None
    if world > 1:
        per = (pop + world - 1) // world
        buf = torch.zeros(per, dtype=torch.float32, device=dev)
        out = torch.empty(per * world, dtype=torch.float32, device=dev)
        for _ in range(5):
            dist.all_gather_into_tensor(out, buf)
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        torch.cuda.synchronize()
        e0.record()
        for _ in range(20):
            dist.all_gather_into_tensor(out, buf)
        e1.record()
        torch.cuda.synchronize()
        gather_us = e0.elapsed_time(e1) / 20 * 1e3
    t_all = torch.tensor([dev_ms, wall], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(t_all, op=dist.ReduceOp.MAX)
    ms_per_step = float(t_all[0].item()) / a.steps
    units = pop * T * em.n_node
    if rank == 0:
        lo, hi = sharding.shard_range(pop, 0, world)
        line = {"metric": METRIC.replace("shunqing GAT rollout", "astlingen MPC population evaluation"), "value": units / (ms_per_step * 1e-3),
                "unit": UNIT, "n_gpus": world, "steps": a.steps, "warmup": max(a.warmup, 3), "ms_per_step": ms_per_step,
                "higher_is_better": True, "scaling": "strong", "vs_baseline": None,
                "dtype": "f32" if precision == "fp32" else ("f16(f32 acc)" if em._lib.gnnuds_tc_operand_is_half() else "bf16(f32 acc)"),
                "data": "synthetic",
                "config": {"workload": "C3mpc: astlingen GAT+Conv1D act+edge_fusion+resnet+flood 5->5, %d candidates TOTAL x %d-step horizon "
                                       "(%d AR chunks, MPC feedback rule) + astlingen objective, %d candidates per GPU" % (pop, T, chunks, hi - lo),
                           "l2": "the chunk loop's activations per pass exceed the 126 MB L2 at %d candidates per GPU" % (hi - lo) if (hi - lo) * T * 59 * 128 * 2 > 126e6 else "small shard: activations of a pass fit the L2 (strong scaling at 8 GPUs)",
                           "precision": precision, "sharding": "candidates over ranks; ONE NCCL all-gather of the objectives per step"},
                "e2e": {"value": units * a.steps / float(t_all[1].item()), "unit": UNIT, "h2d_bytes_per_step": int(x[lo:hi].nbytes + states[0].nbytes + b[0].nbytes + ex[0].nbytes),
                        "d2h_bytes_per_step": int(4 * ((pop + world - 1) // world) * world),
                        "note": "MpcEvaluator.pred_emu_device(x) wall clock: decision matrix in, all-gathered objectives out"},
                "gpu_launches": int(launches), "clocks": clocks, "allgather_us": gather_us,
                "objective_checksum": float(np.sum(F)), "wall_s_timed_region": wall}
        emit(line)
    if world > 1:
        dist.destroy_process_group()


def run_grad(a):
    """--workload C3grad: the gradient MPC's inner call (mpc.py:596-607 gradient_fn) on BASELINE.json configs[2]'s model -
    `--batch` decision vectors (default 256) x 60-step horizon = 12 autoregressive chunks of the astlingen 5->5 model,
    objective AND d objective / d decisions through MpcEvaluator.gradient (gnnuds_grad_settings: taped forward sweep
    and VJP kernels per chunk, fp32; GNNUDS_TAPE_MB=0 = checkpointed sweep).  One GPU; `cpu_baseline` = the gradient oracle (torch fp64 autograd
    through the dense restatement) on a few candidates."""
    import torch
    from gnn_uds_b200 import configs, synthetic, mpc
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if int(os.environ.get("RANK", "0")) != 0:
        return
    torch.cuda.set_device(local)
    em, args, rng = configs.make_emulator("C3b", device=local, precision="fp32")
    net, _ = configs.load_net("astlingen")
    pop = a.batch or 256
    chunks = configs.WORKLOADS["C3b"]["chunks"]
    T = em.seq_out * chunks
    ev = mpc.MpcEvaluator(em, net.nodes, interval=1, eval_horizon=T, control_horizon=T, setting_duration=5,
                          gamma=0.99 ** np.arange(T))
    rng = np.random.default_rng(args.seed)
    states, b, _, ex = synthetic.make_inputs(em, 1, rng, horizon=T)
    ev.load_state(states[0], b[0], ex[0])
    y = mpc.latin_hypercube(pop, np.full(ev.n_var, 0.05), np.ones(ev.n_var), rng)
    for _ in range(max(a.warmup, 3)):
        objs, g = ev.gradient(y)
    sampler = ClockSampler(local)
    sampler.start()
    l0 = em.launch_count()
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    for _ in range(a.steps):
        objs, g = ev.gradient(y)
    torch.cuda.synchronize()
    wall = time.perf_counter() - t0
    launches = em.launch_count() - l0
    clocks = sampler.stop()
    fwd_ms = []
    for _ in range(3):
        t1 = time.perf_counter(); ev.pred_emu(y); fwd_ms.append((time.perf_counter() - t1) * 1e3)
    units = pop * T * em.n_node
    ms = wall / a.steps * 1e3
    # CPU arm: the gradient oracle on a few candidates
    from oracle.emulator_oracle import OracleEmulator
    from oracle import grad_oracle as go
    o = OracleEmulator(args.conv, args.resnet, args.recurrent, args, dtype=np.float64)
    o.set_weights(em.weights)
    o.set_norm(em.norm_x, em.norm_b, em.norm_y, em.norm_r, em.norm_e)
    nc = min(4, pop)
    sett = ev.expand_settings(y[:nc]).astype(np.float64)
    st = np.repeat(states.astype(np.float64), nc, 0); bb = np.repeat(b.astype(np.float64), nc, 0); ee = np.repeat(ex.astype(np.float64), nc, 0)
    gm = 0.99 ** np.arange(T)
    t1 = time.perf_counter()
    ro, rga = go.grad_settings(o, st, bb, sett, ee, ev.fw, ev.ww, ev.rw, gm, chunks=chunks)
    cpu_s = time.perf_counter() - t1
    rel = float(np.abs(ev.collapse_settings_grad(rga) - g[:nc]).max() / max(np.abs(rga).max(), 1e-12))
    emit({"metric": METRIC.replace("shunqing GAT rollout", "astlingen gradient MPC: objective + d/d settings"), "value": units / (ms * 1e-3),
          "unit": UNIT, "n_gpus": 1, "steps": a.steps, "warmup": max(a.warmup, 3), "ms_per_step": ms, "higher_is_better": True,
          "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
          "config": {"workload": "C3grad: astlingen GAT+Conv1D act+edge_fusion+resnet+flood 5->5, %d decision vectors x %d-step horizon (%d AR "
                                 "chunks, MPC feedback), objective and its gradient w.r.t. the %d decisions of every vector" % (pop, T, chunks, ev.n_var),
                     "l2": "tape of a chunk: %.0f MB (%s the 126 MB L2)" % (pop * 4 * 13.0 * 5 * 59 * 128 / 1e6, "exceeds" if pop * 4 * 13.0 * 5 * 59 * 128 > 126e6 else "fits"), "precision": "fp32"},
          "e2e": {"value": units / (ms * 1e-3), "unit": UNIT, "h2d_bytes_per_step": int(pop * T * (em.n_act + em.n_node) * 4),
                  "d2h_bytes_per_step": int(pop * (T * em.n_act + 1) * 4), "note": "MpcEvaluator.gradient(y) wall clock, NumPy in / NumPy out"},
          "gpu_launches": int(launches), "clocks": clocks, "forward_only_ms": float(np.median(fwd_ms)),
          "grad_rel_err_vs_oracle": rel,
          "cpu_baseline": {"value": nc * T * em.n_node / cpu_s, "unit": UNIT, "cores": os.cpu_count(), "kind": "port",
                           "sample": "gradient oracle (torch fp64 autograd, dense formulation) on %d decision vectors, same horizon" % nc}})


def run_c4(a):
    """--workload C4: BASELINE.json configs[3] - the C2 model replayed over all 148 shunqing events (lengths from
    events.csv, synthetic recorded series), whole events balanced over the ranks (strong scaling).  Mode A = the reference's
    test loop (main.py:306-382): every event's T - seq_out teacher-forced windows, built on the device from the raw series
    (Emulator.replay), predictions returned as NumPy arrays.  Mode B (secondary figure) = a true autoregressive roll of
    every event from its first full window (MPC feedback rule)."""
    import torch
    import torch.distributed as dist
    from gnn_uds_b200 import configs, replay, sharding
    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)
    precision = "bf16" if a.precision == "auto" else a.precision
    em, args, _ = configs.make_emulator("C2", device=local, precision=precision)
    names, sets, steps = replay.shunqing_events()
    if a.batch:                                   # --batch K: only the first K events (quick runs)
        names, sets, steps = names[:a.batch], sets[:a.batch], steps[:a.batch]
    mine = sharding.balance_events(steps, world)[rank]
    events = {i: replay.synthetic_event(em, int(steps[i]), np.random.default_rng(1000 + i)) for i in mine}
    To, N = em.seq_out, em.n_node
    units_all = int(sum((int(t) - To) * To * N for t in steps))

    def barrier():
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def one_pass():
        chk, dev_ms = 0.0, 0.0
        if os.environ.get("GNNUDS_C4_PER_EVENT") == "1":            # one Emulator.replay per event (the round-2 start)
            for i in mine:
                e = events[i]
                p, ep = em.replay(e["states"], e["perfs"], e["edge_states"], e["settings"])
                dev_ms += em.last_timing()[0]
                chk += float(p[-1, -1, 0, 0]) + float(ep[-1, -1, 0, 0])
                del p, ep
            return chk, dev_ms
        # Emulator.replay_many: short events share a batch, results of one group leave while the next computes; 32 events
        # per call bound the host memory held at once (~5 GB of predictions)
        order = list(mine)
        for j in range(0, len(order), 32):
            res = em.replay_many([events[i] for i in order[j:j + 32]])
            dev_ms += em.last_timing()[0]
            for p, ep in res:
                chk += float(p[-1, -1, 0, 0]) + float(ep[-1, -1, 0, 0])
            del res
        return chk, dev_ms

    em.set_profiling(1)
    for _ in range(max(1, min(a.warmup, 1))):     # one warm-up pass = 148 events, every shape seen
        one_pass()
    barrier()
    sampler = ClockSampler(local)
    if rank == 0:
        sampler.start()
    l0 = em.launch_count()
    t0 = time.perf_counter()
    dev_ms = 0.0
    for _ in range(a.steps):
        chk, d = one_pass()
        dev_ms += d
    barrier()
    wall = time.perf_counter() - t0
    launches = em.launch_count() - l0
    clocks = sampler.stop() if rank == 0 else None
    # mode B: autoregressive roll of every event of this rank
    tb = time.perf_counter()
    units_b = 0
    for i in mine:
        e = events[i]
        y, _, _ = replay.replay_event_ar(em, e["states"], e["perfs"], e["edge_states"], e["settings"])
        units_b += y.shape[1] * N
    barrier()
    wall_b = time.perf_counter() - tb
    t_all = torch.tensor([dev_ms, wall, wall_b, float(units_b)], dtype=torch.float64, device=dev)
    u_all = torch.tensor([float(units_b)], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(t_all, op=dist.ReduceOp.MAX)
        dist.all_reduce(u_all, op=dist.ReduceOp.SUM)
    if rank == 0:
        ms = float(t_all[0].item()) / a.steps
        line = {"metric": METRIC.replace("rollout", "full-event replay"), "value": units_all / (ms * 1e-3), "unit": UNIT, "n_gpus": world,
                "steps": a.steps, "warmup": 1, "ms_per_step": ms, "higher_is_better": True, "scaling": "strong", "vs_baseline": None,
                "dtype": "f32" if precision == "fp32" else ("f16(f32 acc)" if em._lib.gnnuds_tc_operand_is_half() else "bf16(f32 acc)"),
                "data": "synthetic",
                "config": {"workload": "C4: shunqing GAT+Conv1D edge_fusion resnet flood 60->60 replayed over %d events (%d recorded steps, "
                                       "%d teacher-forced windows, reference accounting sum (T-seq_out)*seq_out*N), one step = all events"
                                       % (len(steps), int(steps.sum()), int(sum(int(t) - To for t in steps))),
                           "l2": "every event's window batch (360..1440 windows x 60 steps) exceeds the 126 MB L2", "precision": precision,
                           "sharding": "whole events balanced over ranks by length, no data-path collective"},
                "e2e": {"value": units_all * a.steps / float(t_all[1].item()), "unit": UNIT,
                        "h2d_bytes_per_step": int(sum(sum(v.nbytes for v in e.values() if v is not None) for e in events.values())),
                        "d2h_bytes_per_step": int(sum((int(steps[i]) - To) * To * (N * em.y_channels + em.n_edge * em.ey_channels) * 4 for i in mine)),
                        "note": "Emulator.replay(raw series) per event, wall clock: raw series in, all window predictions out as NumPy (rank 0's bytes)"},
                "mode_b": {"value": float(u_all.item()) / float(t_all[2].item()), "unit": UNIT,
                           "note": "true autoregressive roll of every event from its first full window (B=1 per event, MPC rule), wall clock"},
                "gpu_launches": int(launches), "clocks": clocks, "checksum": chk, "wall_s_timed_region": wall}
        emit(line)
    if world > 1:
        dist.destroy_process_group()


_JSON_OUT = None


def emit(line):
    """The ONE JSON line goes to the real stdout; everything else (NCCL banners, library chatter
    written to fd 1) has been redirected to stderr by main()."""
    out = _JSON_OUT if _JSON_OUT is not None else sys.stdout
    out.write(json.dumps(line) + "\n")
    out.flush()


def main():
    global _JSON_OUT
    sys.stdout.flush()
    _JSON_OUT = os.fdopen(os.dup(1), "w")
    os.dup2(2, 1)
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=3)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200")
    ap.add_argument("--workload", default="C2")
    ap.add_argument("--batch", type=int, default=0)
    ap.add_argument("--precision", default=os.environ.get("GNNUDS_PRECISION", "auto"))
    ap.add_argument("--no-cpu", action="store_true")
    a = ap.parse_args()
    if a.impl == "reference":
        return run_reference(a)
    if a.workload == "C3mpc":
        return run_mpc(a)
    if a.workload == "C3grad":
        return run_grad(a)
    if a.workload == "C4":
        return run_c4(a)

    import torch
    import torch.distributed as dist
    from gnn_uds_b200 import configs, synthetic, _lib
    import ctypes as C

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py needs a CUDA device (no CPU fallback for the product path)")
    torch.cuda.set_device(local)
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    precision = a.precision
    if precision == "auto":
        precision = "bf16" if os.path.exists(os.path.join(ROOT, "gnn_uds_b200", "csrc", "tc_gemm.cuh")) else "fp32"
    em, args, rng = configs.make_emulator(a.workload, device=local, precision=precision)
    B = a.batch or args.B
    N, M, Ti, To = em.n_node, em.n_edge, em.seq_in, em.seq_out
    # scenarios are independent: every rank owns B of them (weak scaling), no data-path collective
    rng = np.random.default_rng(args.seed + 1000 * rank)
    Bg = min(B, 256)
    states, b, act, ex = synthetic.make_inputs(em, Bg, rng)
    rep = (B + Bg - 1) // Bg
    tile = lambda t: None if t is None else np.ascontiguousarray(np.concatenate([t] * rep, 0)[:B])
    states, b, act, ex = tile(states), tile(b), tile(act), tile(ex)
    yc, ec = em.y_channels, em.ey_channels

    # ---- device-resident arm -------------------------------------------------
    dev = torch.device("cuda", local)
    dx, db, dex = [torch.from_numpy(t).to(dev) for t in (states, b, ex)]
    da = torch.from_numpy(act).to(dev) if act is not None else None
    dy = torch.empty((B, To, N, yc), dtype=torch.float32, device=dev)
    dey = torch.empty((B, To, M, ec), dtype=torch.float32, device=dev)
    torch.cuda.synchronize()
    in_bytes = sum(t.numel() * 4 for t in (dx, db, dex)) + (da.numel() * 4 if da is not None else 0)
    l2_note = "inputs %.0f MB + activations per pass exceed the 126 MB L2" % (in_bytes / 1e6)

    def step_dev():
        em.predict_ptrs(dx.data_ptr(), db.data_ptr(), da.data_ptr() if da is not None else None, dex.data_ptr(), B,
                        dy.data_ptr(), dey.data_ptr(), device=True)

    def barrier():
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    for _ in range(max(a.warmup, 3)):
        step_dev()
    barrier()
    sampler = ClockSampler(local)
    if rank == 0:
        sampler.start()
    l0 = em.launch_count()
    em.set_profiling(2)             # CUDA events on the library's own stream (torch events cannot see it); 2 = also per-GEMM pairs
    tot_ms, gemm_ms = 0.0, 0.0
    dom = dominant_kernel(em, B, precision)
    dom_ms, dom_launches = 0.0, 0
    fam_ms = [0.0] * 4
    for _ in range(a.steps):
        step_dev()
        t, g = em.last_timing()
        tot_ms += t; gemm_ms += g
        fm, fl = em.kernel_timing(dom["family"])
        dom_ms += fm; dom_launches += fl
        for f in range(4):
            fam_ms[f] += em.kernel_timing(f)[0]
    barrier()
    launches = em.launch_count() - l0
    # the per-GEMM event pairs perturb the stream slightly; time the K steps again without them
    em.set_profiling(False)
    barrier()
    t0 = time.perf_counter()
    ev_ms = 0.0
    em.set_profiling(1)             # whole-step event pair only
    for _ in range(a.steps):
        step_dev()
        ev_ms += em.last_timing()[0]
    barrier()
    wall = time.perf_counter() - t0
    clocks = sampler.stop() if rank == 0 else None
    dev_ms = ev_ms                         # device time of exactly K steps (CUDA events, library stream)
    t_all = torch.tensor([dev_ms], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(t_all, op=dist.ReduceOp.MAX)
    ms_per_step = float(t_all.item()) / a.steps
    value = world * B * To * N / (ms_per_step * 1e-3)

    # ---- end-to-end arm: host (pinned) buffers through Emulator.predict ---------
    def pinned(shape):
        n = int(np.prod(shape))
        p = C.c_void_p()
        rc = em._lib.gnnuds_alloc_pinned(C.byref(p), n * 4)
        assert rc == 0
        return np.ctypeslib.as_array(C.cast(p, C.POINTER(C.c_float)), shape=(n,)).reshape(shape), p
    hx, px = pinned(states.shape); hx[...] = states
    hb, pb = pinned(b.shape); hb[...] = b
    hex_, pex = pinned(ex.shape); hex_[...] = ex
    ha, pa = (None, None)
    if act is not None:
        ha, pa = pinned(act.shape); ha[...] = act
    hy, py = pinned((B, To, N, yc)); hey, pey = pinned((B, To, M, ec))

    def step_e2e():
        em.predict_ptrs(px, pb, pa, pex, B, py, pey, device=False)
        return float(hy[0, 0, 0, 0])

    em.set_profiling(False)
    for _ in range(2):
        step_e2e()
    barrier()
    t0 = time.perf_counter()
    for _ in range(a.steps):
        step_e2e()
    barrier()
    e2e_s = time.perf_counter() - t0
    e_all = torch.tensor([e2e_s], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(e_all, op=dist.ReduceOp.MAX)
    e2e_val = world * B * To * N * a.steps / float(e_all.item())
    h2d = int(in_bytes)
    d2h = int((dy.numel() + dey.numel()) * 4)
    # the caller's real entry: Emulator.predict on ordinary (pageable) NumPy arrays, results as NumPy arrays
    res = None
    for _ in range(2):
        res = em.predict(states, b, act, ex)
    del res
    barrier()
    t0 = time.perf_counter()
    chk = 0.0
    for _ in range(a.steps):
        res = em.predict(states, b, act, ex)
        chk += float(res[0][0, 0, 0, 0])
        del res
    barrier()
    np_s = time.perf_counter() - t0
    n_all = torch.tensor([np_s], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(n_all, op=dist.ReduceOp.MAX)
    np_val = world * B * To * N * a.steps / float(n_all.item())

    if rank == 0:
        tot_f, gemm_f = flops(em, B)
        tf_peak, hbm_peak, src = peaks()
        gemm_ms_per_step = gemm_ms / a.steps
        dom_ms_per_step = dom_ms / a.steps
        launches_per_step = max(dom_launches // a.steps, 1)
        extra = {"kernel": dom["kernel"], "launches_per_step": launches_per_step, "kernel_ms_per_step": dom_ms_per_step,
                 "avg_launch_ms": dom_ms_per_step / launches_per_step, "share_of_step": dom_ms_per_step / ms_per_step,
                 "algorithmic": dom["unit_note"], "total_flops_per_step": tot_f,
                 "step_tflops": tot_f / (ms_per_step * 1e-3) / 1e12, "step_tflops_frac_of_bf16_peak": tot_f / (ms_per_step * 1e-3) / 1e12 / tf_peak,
                 "family_ms_per_step": {"rows_gemm": fam_ms[0] / a.steps, "fused_spatial": fam_ms[1] / a.steps,
                                        "csr_gather": fam_ms[2] / a.steps, "recurrent": fam_ms[3] / a.steps,
                                        "note": "per-launch event pairs of a separate K-step pass; the rest of the step is embed/heads/post-processing"}}
        if dom["bound"] == "tensor" and dom.get("flops"):
            # C1-C4 are tensor-bound; achieved = FLOPs the fused spatial kernel executes per launch / launch time
            fpl = dom["flops"] / launches_per_step
            avg_ms = dom_ms_per_step / launches_per_step
            achieved = fpl / (avg_ms * 1e-3) / 1e12 if avg_ms > 0 else None
            hbm_ach = dom["bytes"] / (dom_ms_per_step * 1e-3) / 1e9 if dom_ms_per_step > 0 else None
            tr = traffic_per_launch(dom, a.workload, B, launches_per_step)
            roofline = {"bound": "tensor", "achieved": achieved, "peak": tf_peak, "unit": "TFLOP/s",
                        "frac": (achieved / tf_peak) if achieved else None, "traffic": tr,
                        "algorithmic_flops_per_launch": fpl, "peak_source": "%s bf16_tflops_sustained" % src,
                        "design_hbm_frac": (hbm_ach / hbm_peak) if hbm_ach else None,
                        "design_hbm_bytes_per_launch": dom["bytes"] / launches_per_step,
                        "io_bytes_per_step": dom["io_bytes"],
                        "traffic_over_io": (tr * launches_per_step / dom["io_bytes"]) if tr else None}
        elif dom["bound"] == "fp32":
            achieved = gemm_f / (gemm_ms_per_step * 1e-3) / 1e12 if gemm_ms_per_step > 0 else None
            f32_peak = em.measure_fp32_tflops()
            roofline = {"bound": "tensor", "achieved": achieved, "peak": f32_peak, "unit": "TFLOP/s",
                        "frac": (achieved / f32_peak) if achieved and f32_peak else None, "traffic": None,
                        "peak_source": "fp32 FFMA peak measured in this run (gnnuds_measure_simt_tflops); the fp32 parity path runs on SIMT FMA, not tensor cores",
                        "gemm_flops_per_step": gemm_f}
        elif dom["bound"] == "hbm":
            achieved = dom["bytes"] / (dom_ms_per_step * 1e-3) / 1e9 if dom_ms_per_step > 0 else None
            roofline = {"bound": "hbm", "achieved": achieved, "peak": hbm_peak, "unit": "GB/s",
                        "frac": (achieved / hbm_peak) if achieved else None,
                        "traffic": traffic_per_launch(dom, a.workload, B, launches_per_step),
                        "algorithmic_bytes_per_launch": dom["bytes"] / launches_per_step, "peak_source": "%s hbm_gbs" % src}
        else:
            achieved = gemm_f / (gemm_ms_per_step * 1e-3) / 1e12 if gemm_ms_per_step > 0 else None
            roofline = {"bound": "tensor", "achieved": achieved, "peak": tf_peak, "unit": "TFLOP/s",
                        "frac": (achieved / tf_peak) if achieved else None, "traffic": None,
                        "peak_source": "%s bf16_tflops_sustained" % src, "gemm_flops_per_step": gemm_f}
        roofline.update(extra)
        line = {
            "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": a.steps, "warmup": max(a.warmup, 3),
            "ms_per_step": ms_per_step, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
            "dtype": "f32" if (precision == "fp32" or em.conv_kind == "nn") else ("f16(f32 acc)" if em._lib.gnnuds_tc_operand_is_half() else "bf16(f32 acc)"), "data": "synthetic",
            "config": {"workload": "%s: %s %s+%s edge_fusion=%s resnet=%s if_flood=%d act=%s, N=%d M=%d, seq %d->%d, B=%d per GPU" % (
                a.workload, args.__dict__.get("env_name", configs.WORKLOADS[a.workload]["net"]), em.conv_kind.upper(), em.recurrent,
                em.edge_fusion, em.resnet, em.if_flood, bool(em.act), N, M, Ti, To, B),
                "l2": l2_note, "precision": ("f16" if em._lib.gnnuds_tc_operand_is_half() else "bf16") if precision == "bf16" else precision, "sharding": "scenarios over ranks, no data-path collective"},
            "e2e": {"value": e2e_val, "unit": UNIT, "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": d2h,
                    "note": "Emulator.predict_ptrs with pinned host buffers, wall clock incl. copies",
                    "numpy_api": {"value": np_val, "unit": UNIT, "frac_of_device_resident": np_val / value,
                                  "note": "Emulator.predict(states, b, a, edge_state) on pageable NumPy arrays, NumPy results; "
                                          "the library stages chunks through a pinned ring with copy threads"}},
            "gpu_launches": int(launches),
            "clocks": clocks,
            "roofline": roofline,
            "wall_s_timed_region": wall,
        }
        if world == 1 and not a.no_cpu:
            line["cpu_baseline"] = cpu_baseline(a.workload)
            # "C1 exactly as BASELINE.json words it" (astlingen GCN+GRU 10->10, B=64), median of 5
            line["cpu_baseline_c1"] = cpu_baseline("C1", fixed_B=64, runs=5)
        emit(line)
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
