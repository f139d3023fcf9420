#!/usr/bin/env python
"""bench.py -- train graphs/s of the GraphCNNDropEdge hot path on N B200s (BASELINE.json metric).

  python bench.py --gpus 1 --steps K --warmup W              our arm (CUDA kernels behind the C ABI)
  python bench.py --impl reference --gpus 1 --steps K ...    the reference's CPU implementation (oracle port)
  torchrun --nproc-per-node N bench.py --gpus N ...          data parallel, one process per GPU (weak scaling)

A "step" is one KVProcedure train step (kv_procedure.py:143-164: forward([V, A]) -> CE -> backward -> clip -> Adam)
over one synthetic batch of the named workload.  One JSON line on stdout (rank 0).
"""
from __future__ import annotations

import argparse
import json
import os
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

METRIC = "train graphs/sec (fwd+bwd)"
UNIT = "graphs/s"


def parse_args():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--workload", default="cfg2", choices=["cfg1", "cfg2", "cfg3"])
    ap.add_argument("--batch", type=int, default=None, help="override graphs per GPU (debug)")
    ap.add_argument("--cpu-batch", type=int, default=8, help="graphs per step of the bounded CPU sample")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-e2e", action="store_true")
    ap.add_argument("--no-roofline", action="store_true")
    ap.add_argument("--no-bow", action="store_true", help="skip the bag-of-words feature variant (sparse emb1 path)")
    ap.add_argument("--x3-promo", type=int, default=None, help="K-blocks per TMEM accumulator promotion of the fp32 forward GEMMs (A/B)")
    ap.add_argument("--profile-step", action="store_true",
                    help="bracket ONE extra train step (after the timed region) with cudaProfilerStart/Stop for "
                         "`ncu --profile-from-start off`; numbers printed by such a run are not bench values")
    return ap.parse_args()


# ------------------------------------------------------------------------------------------------ helpers
def peaks():
    path = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(path):
        p = json.load(open(path))
        return {"hbm_gbs": p["hbm_gbs"], "bf16_tflops": p["bf16_tflops"],
                "bf16_tflops_sustained": p.get("bf16_tflops_sustained", p["bf16_tflops"]), "source": "measured"}
    return {"hbm_gbs": 6650.0, "bf16_tflops": 1590.0, "bf16_tflops_sustained": 1400.0, "source": "fallback"}


class ClockSampler(threading.Thread):
    """nvidia-smi clocks / throttle reasons sampled DURING the timed region (B200_PROFILING.md recipe)."""
    Q = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
         "clocks_event_reasons.sw_power_cap")

    def __init__(self, index: int):
        super().__init__(daemon=True)
        self.index, self.rows, self._halt = index, [], threading.Event()

    def run(self):
        while not self._halt.is_set():
            try:
                out = subprocess.run(["nvidia-smi", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits",
                                      "-i", str(self.index)], capture_output=True, text=True, timeout=5).stdout
                parts = [x.strip() for x in out.strip().split(",")]
                if len(parts) >= 7:
                    self.rows.append(parts)
            except Exception:
                pass
            self._halt.wait(0.2)

    def finish(self):
        self._halt.set()
        self.join(timeout=3)
        if not self.rows:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["unsampled"]}
        sm = sorted(float(r[0]) for r in self.rows)
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        reasons = [n for k, n in enumerate(names) if any(r[3 + k].lower().startswith("active") for r in self.rows)]
        return {"sm_mhz": sm[len(sm) // 2], "sm_max_mhz": float(self.rows[0][1]), "reasons": reasons,
                "samples": len(self.rows)}


def cpu_reference_arm(wl, B, steps, warmup):
    """The reference's CPU path (oracle port of drop_robust_gcn.py:61-103 + kv_procedure.py:143-164) on the host
    cores, dense adjacency, nn.Dropout semantics.  Returns (graphs/s, ms/step, cores)."""
    import torch
    from graph_representation_learning_b200 import GraphCNNDropEdge, synthetic
    from oracle import reference_port as rp
    cores = os.cpu_count() or 1
    torch.set_num_threads(cores)
    torch.manual_seed(0)
    model = GraphCNNDropEdge(wl.input_dim, wl.output_dim, wl.num_edges, wl.net_size, wl.use_attention)
    sd = {k: v.detach().clone() for k, v in model.state_dict().items()}
    V, A, y = synthetic.make_batch(wl, 1234, "cpu", B=B)
    A = A.float()                                            # kv_procedure.py:118
    state = {}
    times = []
    for it in range(warmup + steps):
        t0 = time.perf_counter()
        rp.train_step(sd, V, A, y, state, step=it + 1)
        times.append(time.perf_counter() - t0)
    t = sum(times[warmup:]) / steps
    return B / t, t * 1e3, cores


# ------------------------------------------------------------------------------------------------ reference arm
def run_reference(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    from graph_representation_learning_b200 import synthetic
    wl = synthetic.CONFIGS[args.workload]
    B = args.cpu_batch
    steps, warmup = max(1, min(args.steps, 5)), max(1, min(args.warmup, 2))
    val, ms, cores = cpu_reference_arm(wl, B, steps, warmup)
    sample = f"{B} graphs x {wl.N} nodes per step ({warmup}+{steps} steps) of {wl.name}, dense fp32 adjacency"
    print(json.dumps({
        "impl": "reference", "metric": METRIC, "value": val, "unit": UNIT, "n_gpus": args.gpus, "steps": steps,
        "warmup": warmup, "ms_per_step": ms, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
        "dtype": "f32", "data": "synthetic",
        "config": {"workload": wl.name, "step": "fwd+bwd+clip+adam", "graphs_per_step": B},
        "cpu_baseline": {"value": val, "unit": UNIT, "cores": cores, "kind": "port", "sample": sample},
        "e2e": {"value": val, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
    }))


# ------------------------------------------------------------------------------------------------ our arm
def run_b200(args):
    import torch
    import torch.distributed as dist
    from graph_representation_learning_b200 import GraphCNNDropEdge, PackedGraphBatch, _lib, ops, synthetic
    from graph_representation_learning_b200.trainer import KVTrainStep

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    if world != args.gpus:
        raise SystemExit(f"--gpus {args.gpus} but WORLD_SIZE={world}: launch with torchrun --nproc-per-node {args.gpus}")
    dev = torch.device("cuda", local_rank)
    numa_cpus = None
    if world > 1:       # host buffers and copy threads next to this rank's GPU (before any pinned allocation)
        from graph_representation_learning_b200.trainer import bind_process_to_gpu_numa
        numa_cpus = bind_process_to_gpu_numa(local_rank)
    torch.cuda.set_device(dev)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)
    if _lib.load().grl_device_ok() != 1:
        raise SystemExit("bench.py needs a compute-capability 10.x GPU (B200); there is no fallback path")

    wl = synthetic.CONFIGS[args.workload]
    B = args.batch or wl.B
    if args.x3_promo:
        ops.set_x3_promotion(args.x3_promo)
    torch.manual_seed(0)                                     # identical replicas
    model = GraphCNNDropEdge(wl.input_dim, wl.output_dim, wl.num_edges, wl.net_size, wl.use_attention).to(dev)
    model.set_compute_dtype(wl.dtype)
    model.rng_seed = 1000 + rank                             # independent DropEdge/dropout streams per rank
    step = KVTrainStep(model)
    feat_dtype = torch.float32 if wl.dtype == torch.float32 else torch.bfloat16
    V, A, y = synthetic.make_batch(wl, 1234 + rank, dev, B=B, feature_dtype=feat_dtype)
    batch_dev = {"textline_encoding": V, "adjacency_matrix": A, "node_label": y}

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def timed(fn, steps, warmup):
        for _ in range(warmup):
            fn()
        barrier()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        n0 = _lib.launches()
        sampler = ClockSampler(local_rank) if rank == 0 else None
        if sampler:
            sampler.start()
        e0.record()
        for _ in range(steps):
            fn()
        e1.record()
        barrier()
        ms = e0.elapsed_time(e1)
        clocks = sampler.finish() if sampler else None
        t = torch.tensor([ms], device=dev)
        if world > 1:
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t) / steps, _lib.launches() - n0, clocks

    # (1) device-resident inputs: the whole forward([V, A]) API call incl. the CSR/CSC pack of the dense fp16 A
    ms_step, launches, clocks = timed(lambda: step.run_train_step(batch_dev), args.steps, args.warmup)
    value = world * B / (ms_step * 1e-3)

    if args.profile_step:
        torch.cuda.synchronize()
        torch.cuda.profiler.start()
        step.run_train_step(batch_dev)
        torch.cuda.synchronize()
        torch.cuda.profiler.stop()

    # (2) end to end: pinned HOST buffers -> H2D inside the timed region, loss read back to the host.  Two host loaders of
    # the SAME batch: the EdgeListCollate format (features + (b, i, l, j) triples; the adjacency never exists densely,
    # SURVEY 8f N1) is the headline; the reference's dense fp16 [B, N, L, N] adjacency is reported beside it.
    e2e = None
    if not args.no_e2e:
        from graph_representation_learning_b200.trainer import DevicePrefetcher

        def pin(t):
            return torch.empty(t.shape, dtype=t.dtype, pin_memory=True).copy_(t)

        Vp = V._base if V._base is not None else V          # the padded storage (16-byte row pitch)
        host_dense = {k: pin(v) for k, v in batch_dev.items()}
        Vp_h = pin(Vp)
        host_edges = {"textline_encoding": Vp_h[..., :wl.input_dim], "features_padded": Vp_h,
                      "node_label": host_dense["node_label"], "edges": pin(A.nonzero().to(torch.int32)),
                      "edge_weights": None, "graph_shape": (B, wl.N, wl.num_edges)}
        pinned_loss = torch.zeros(64, dtype=torch.float32).pin_memory()

        def nbytes(d, keys):
            return sum(d[k].numel() * d[k].element_size() for k in keys)

        def e2e_run(host, steps, warmup):
            """the public step API over a host loader: H2D of every step's inputs (pinned -> device, one batch ahead on
            a copy stream) and a D2H read of every step's loss, all inside the timed region"""
            barrier()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            seen = []
            for i, dev_batch in enumerate(DevicePrefetcher((host for _ in range(warmup + steps)), dev)):
                if i == warmup:
                    e0.record()
                loss = step.run_train_step(dev_batch)
                pinned_loss[i % 64].copy_(loss, non_blocking=True)
                ev = torch.cuda.Event()
                ev.record()
                if seen:                                   # read the previous step's loss on the host
                    pev, pi = seen[-1]
                    pev.synchronize()
                    float(pinned_loss[pi % 64])
                seen.append((ev, i))
            e1.record()
            barrier()
            t = torch.tensor([e0.elapsed_time(e1)], device=dev)
            if world > 1:
                dist.all_reduce(t, op=dist.ReduceOp.MAX)
            return float(t) / steps
        e_steps = max(3, min(args.steps, 10))
        ms_edges = e2e_run(host_edges, e_steps, 2)
        ms_dense = e2e_run(host_dense, e_steps, 2)
        e2e = {"value": world * B / (ms_edges * 1e-3), "unit": UNIT,
               "h2d_bytes_per_step": nbytes(host_edges, ["features_padded", "node_label", "edges"]),
               "d2h_bytes_per_step": 4, "ms_per_step": ms_edges, "steps": e_steps,
               "how": "KVTrainStep.run_train_step over DevicePrefetcher(host loader of EdgeListCollate batches): pinned "
                      "host V (fp32, all columns) / y / int32 (b,i,l,j) edge triples -> device one batch ahead on a copy "
                      "stream, CSR/CSC packed on the device, loss read back every step",
               "dense_adjacency": {"value": world * B / (ms_dense * 1e-3), "ms_per_step": ms_dense,
                                   "h2d_bytes_per_step": nbytes(host_dense, list(host_dense)),
                                   "how": "same, host loader yields the reference's dense fp16 [B,N,L,N] adjacency "
                                          "(NumpyPadding + default_collate format)"}}
        del host_dense, host_edges, Vp_h

    # (2b) the second named feature variant: bag-of-words text-line encodings (7 ones in the 4365 binary
    # columns + 4 spatial floats per node, what TextlineEncoding emits) through the sparse-feature path -- same model,
    # same graph, same step; emb1 runs as a gather-sum + a 4-column GEMM (features.py), the host ships CSR + edge triples
    bow = None
    if not args.no_bow and wl.dtype == torch.float32:
        from graph_representation_learning_b200 import SparseNodeFeatures
        from graph_representation_learning_b200.trainer import DevicePrefetcher
        Vb, _, _ = synthetic.make_batch(wl, 1234 + rank, dev, B=B, features="bow")
        sf = SparseNodeFeatures.from_dense(Vb)
        del Vb

        def fresh():   # a new batch object every step: the CSC of the features is rebuilt inside the step, like A's
            return SparseNodeFeatures(sf.shape, sf.rowptr, sf.col, sf.val, sf.heavy_cols, sf.heavy)
        ms_bow, _, _ = timed(lambda: step.run_train_step({"textline_encoding": fresh(), "adjacency_matrix": A,
                                                          "node_label": y}), max(5, args.steps // 2), 3)
        bow = {"value": world * B / (ms_bow * 1e-3), "unit": UNIT, "ms_per_step": ms_bow,
               "features": "bow: 7 ones in 4365 binary columns + 4 floats per node, shipped as CSR (%d entries) + a "
                           "[Nt, %d] heavy-column block" % (sf.nnz, int(sf.heavy_cols.numel()))}
        if not args.no_e2e:
            host_bow = {"textline_encoding": sf.to("cpu").pin_memory(), "node_label": pin(y),
                        "edges": pin(A.nonzero().to(torch.int32)), "edge_weights": None,
                        "graph_shape": (B, wl.N, wl.num_edges)}
            ms_bow_e2e = e2e_run(host_bow, e_steps, 2)
            bow["e2e"] = {"value": world * B / (ms_bow_e2e * 1e-3), "unit": UNIT, "ms_per_step": ms_bow_e2e,
                          "h2d_bytes_per_step": host_bow["textline_encoding"].nbytes() + nbytes(host_bow, ["node_label", "edges"]),
                          "d2h_bytes_per_step": 4}
            del host_bow
        del sf

    # (3) roofline of the aggregation kernel (gcn1 SpMM, DropEdge weights applied), timed alone with CUDA events
    pk = peaks()
    g = PackedGraphBatch.from_dense(A)
    S = wl.net_size
    X = torch.randn(B * wl.N, S, device=dev, dtype=wl.dtype)
    Y = torch.empty(g.num_rows, S, device=dev, dtype=wl.dtype)
    w = g.edge_weights(0.3, seed=1, stream_id=1)
    ms_spmm, _, _ = timed(lambda: ops.spmm_fwd_(g, w, X, Y), 20, 5)
    es = X.element_size()
    alg_bytes = 4 * (g.num_rows + 1) + 8 * g.nnz + X.numel() * es + Y.numel() * es   # SURVEY 8d bytes_compulsory (+w)
    achieved = alg_bytes / (ms_spmm * 1e-3) / 1e9
    traffic, traffic_bwd = None, None               # ncu dram__bytes_read+write per launch (profiles/, cfg2 shape only)
    tpath = os.path.join(ROOT, "profiles", "spmm_traffic.json")
    if wl.name.startswith("cfg2") and B == wl.B and os.path.exists(tpath):
        tj = json.load(open(tpath))
        traffic, traffic_bwd = tj["fwd"]["dram_bytes_per_launch"], tj["bwd"]["dram_bytes_per_launch"]
    roofline = {"kernel": "spmm_slab2_kernel (gcn1 aggregation fwd over the DropEdge node streams)", "bound": "hbm",
                "achieved": achieved, "peak": pk["hbm_gbs"], "unit": "GB/s", "frac": achieved / pk["hbm_gbs"],
                "traffic": traffic,
                "peak_source": pk["source"], "ms_per_launch": ms_spmm, "algorithmic_bytes": alg_bytes}
    dYt = torch.randn(g.num_rows, S, device=dev, dtype=wl.dtype)
    dX = torch.empty(B * wl.N, S, device=dev, dtype=wl.dtype)
    ms_bwd, _, _ = timed(lambda: ops.spmm_bwd_(g, w, dYt, dX), 20, 5)
    alg_bwd = 4 * (g.num_nodes + 1) + 12 * g.nnz + dYt.numel() * es + dX.numel() * es
    roofline["bwd"] = {"kernel": "spmm_rows_kernel<transposed>", "achieved": alg_bwd / (ms_bwd * 1e-3) / 1e9,
                       "frac": alg_bwd / (ms_bwd * 1e-3) / 1e9 / pk["hbm_gbs"], "ms_per_launch": ms_bwd,
                       "algorithmic_bytes": alg_bwd, "traffic": traffic_bwd}
    del X, Y, dYt, dX
    # GEMMs are the larger share of the step (about 60 %): the gcn3 projection forward [Nt x S, K = 7 * 2S] and its dX,
    # timed alone.  Tensor-bound figure = EXECUTED tensor flops (fp32: three tf32 products forward, tf32 + double-K bf16
    # backward = 2 tf32-product equivalents; bf16: one product) against the dense peak of that operand type
    # (tf32 = half the measured bf16 figure).  "useful" = 2*M*N*K of the fp32-grade product.
    Kg = 14 * S
    Xg = torch.randn(B * wl.N, Kg, device=dev, dtype=wl.dtype)
    Wg = torch.randn(S, Kg, device=dev) / Kg ** 0.5
    dZg = torch.randn(B * wl.N, S, device=dev, dtype=wl.dtype)
    ms_gf, _, _ = timed(lambda: ops.gemm_nt([Xg], [Wg]), 10, 3)
    ms_gb, _, _ = timed(lambda: ops.gemm_nt([dZg], [Wg.t()], fp32_mode=ops.GemmConfig.fp32_mode_bwd), 10, 3)
    useful = 2.0 * B * wl.N * S * Kg
    fp32 = wl.dtype == torch.float32
    peak_tf = pk["bf16_tflops_sustained"] / 2 if fp32 else pk["bf16_tflops_sustained"]
    prod_f, prod_b = (3.0, 2.0) if fp32 else (1.0, 1.0)
    roofline["gemm"] = {
        "bound": "tensor", "unit": "TFLOP/s", "peak": peak_tf,
        "peak_how": "measured sustained cuBLAS bf16" + (" / 2 (tf32 operands)" if fp32 else ""),
        "fwd": {"kernel": "gemm_x3_kernel (gcn3 projection, three tf32 products)" if fp32 else "gemm_nt_bf16_persistent_kernel (gcn3 projection)",
                "shape_mnk": [B * wl.N, S, Kg], "ms_per_launch": ms_gf, "useful": useful / ms_gf / 1e9,
                "achieved": prod_f * useful / ms_gf / 1e9, "frac": prod_f * useful / ms_gf / 1e9 / peak_tf},
        "bwd_dx": {"kernel": "gemm_bc_kernel<NT> (gcn3 dX, tf32 + bf16 correction product)" if fp32 else "gemm_nt_bf16_persistent_kernel (gcn3 dX)",
                   "shape_mnk": [B * wl.N, Kg, S], "ms_per_launch": ms_gb, "useful": useful / ms_gb / 1e9,
                   "achieved": prod_b * useful / ms_gb / 1e9, "frac": prod_b * useful / ms_gb / 1e9 / peak_tf},
        "note": "all GEMMs of the step run under the board power cap (sw_power_cap, ~1000 W) and are L2->SM bound "
                "(DESIGN.md 4.2); the fractions are against an un-capped cuBLAS figure"}
    del Xg, Wg, dZg

    # (4) CPU baseline: the oracle port on a bounded sample of the same workload (rank 0, N=1 only)
    cpu = None
    if rank == 0 and world == 1 and not args.no_cpu_baseline:
        cb = args.cpu_batch
        v, ms_cpu, cores = cpu_reference_arm(wl, cb, 3, 1)
        cpu = {"value": v, "unit": UNIT, "cores": cores, "kind": "port", "ms_per_step": ms_cpu,
               "sample": f"{cb} graphs x {wl.N} nodes per step (1+3 steps) of the same workload, dense fp32 adjacency"}

    if rank == 0:
        print(json.dumps({
            "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
            "ms_per_step": ms_step, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
            "dtype": "f32" if wl.dtype == torch.float32 else "bf16", "data": "synthetic",
            "config": {"workload": wl.name, "step": "fwd+bwd+clip+adam (forward includes the CSR/CSC pack of dense A)",
                       "graphs_per_gpu": B, "global_batch": world * B, "parallelism": f"dp{world}",
                       "l2": "inputs (V %.2f GB) larger than L2" % (V.numel() * V.element_size() / 1e9),
                       "features": "dense U[0,1)",
                       "host_cpus_rank0": (f"{len(numa_cpus)} GPU-local cores" if numa_cpus else "unbound")},
            "e2e": e2e, "gpu_launches": launches, "clocks": clocks, "roofline": roofline, "cpu_baseline": cpu,
            "bow_features": bow,
        }))
    if world > 1:
        dist.destroy_process_group()


def main():
    args = parse_args()
    # stdout carries exactly ONE JSON line: anything native libraries print there (NCCL's version banner under
    # NCCL_DEBUG=VERSION, ...) is sent to stderr by pointing fd 1 at fd 2 while the run lasts
    sys.stdout.flush()
    real_stdout = os.dup(1)
    os.dup2(2, 1)
    try:
        import contextlib
        import io
        buf = io.StringIO()
        with contextlib.redirect_stdout(buf):
            if args.impl == "reference":
                run_reference(args)
            else:
                run_b200(args)
    finally:
        sys.stdout.flush()
        os.dup2(real_stdout, 1)
        os.close(real_stdout)
    lines = [ln for ln in buf.getvalue().splitlines() if ln.startswith("{")]
    if lines:
        print(lines[-1], flush=True)


if __name__ == "__main__":
    main()
