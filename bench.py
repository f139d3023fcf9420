#!/usr/bin/env python
"""bench.py -- train graphs/s of the GraphCNNDropEdge hot path on N B200s (BASELINE.json metric).

  python bench.py --gpus 1 --steps K --warmup W              our arm (CUDA kernels behind the C ABI)
  python bench.py --impl reference --gpus 1 --steps K ...    the reference's own CPU implementation (oracle/_ref)
  torchrun --nproc-per-node N bench.py --gpus N ...          data parallel, one process per GPU (weak scaling)

A "step" is one KVProcedure train step (kv_procedure.py:143-164: forward([V, A]) -> CE -> backward -> clip -> Adam)
over one synthetic batch of the named workload.  One JSON line on stdout (rank 0).

Line layout: the contract keys are about cfg2 (BASELINE.json configs[1]: 256 graphs x 500 nodes, S=256, fp32); the other
BASELINE configurations ride in objects of the same line -- `cfg3` (S=512 bf16, 1000-node graphs: weak at 128 graphs / GPU
and strong at a global batch of 1024), and at N = 1 `cfg4` (one 1M-node graph) and `cfg5` (eval latency sweep).
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
NOMINAL_HBM_GBS = 8000.0          # the figure BASELINE.json's north_star quotes ("about 8 TB/s")


def parse_args():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--workload", default="cfg2", choices=["cfg1", "cfg2", "cfg3"])
    ap.add_argument("--batch", type=int, default=None, help="override graphs per GPU (debug)")
    ap.add_argument("--cpu-batch", type=int, default=16, help="graphs per step of the bounded CPU sample (SURVEY 8d: 16)")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-e2e", action="store_true")
    ap.add_argument("--no-roofline", action="store_true")
    ap.add_argument("--no-bow", action="store_true", help="skip the bag-of-words feature variant (sparse emb1 path)")
    ap.add_argument("--no-configs", action="store_true", help="skip the cfg3 / cfg4 / cfg5 objects")
    ap.add_argument("--x3-promo", type=int, default=None, help="K-blocks per TMEM accumulator promotion of the fp32 forward GEMMs (A/B)")
    ap.add_argument("--profile-step", action="store_true",
                    help="bracket ONE extra train step (after the timed region) with cudaProfilerStart/Stop for "
                         "`ncu --profile-from-start off`; numbers printed by such a run are not bench values")
    ap.add_argument("--profile-workload", default=None, choices=[None, "cfg3", "cfg4"],
                    help="with --profile-step: profile one step of this configuration instead of the bench line's")
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
            self._halt.wait(0.1)

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


# ------------------------------------------------------------------------------------------------ CPU arm
class CpuReference:
    """The reference's CPU path on the host cores: the UNMODIFIED ``gnn.models.GraphCNNDropEdge`` staged under
    oracle/_ref (oracle/ref_loader.py) inside a restated ``_run_train_step`` (kv_procedure.py:143-164), dense fp32
    adjacency and nn.Dropout exactly like the reference.  Falls back to the port (oracle/reference_port.py) only when
    the staged reference is absent, and says so in ``kind``."""

    def __init__(self, wl, threads=None):
        import torch
        from oracle import ref_loader
        self.torch, self.wl = torch, wl
        self.cores = threads or os.cpu_count() or 1
        torch.set_num_threads(self.cores)
        loaded = ref_loader.load_reference()
        torch.manual_seed(0)
        if loaded is not None:
            self.kind = "reference"
            self.model = loaded[0](wl.input_dim, wl.output_dim, wl.num_edges, net_size=wl.net_size,
                                   use_attention=wl.use_attention)
            self.opt = torch.optim.Adam([p for p in self.model.parameters() if p.requires_grad])
            self._step = lambda V, A, y: ref_loader.reference_train_step(self.model, self.opt, V, A, y)
        else:
            from graph_representation_learning_b200 import GraphCNNDropEdge
            from oracle import reference_port as rp
            self.kind = "port"
            self.model = GraphCNNDropEdge(wl.input_dim, wl.output_dim, wl.num_edges, wl.net_size, wl.use_attention)
            sd = {k: v.detach().clone() for k, v in self.model.state_dict().items()}
            self._sd = sd
            state, it = {}, [0]

            def step(V, A, y):
                it[0] += 1
                return rp.train_step(sd, V, A, y, state, step=it[0])[1]
            self._step = step

    def batch(self, B):
        from graph_representation_learning_b200 import synthetic
        V, A, y = synthetic.make_batch(self.wl, 1234, "cpu", B=B)
        return V.contiguous(), A.float(), y                  # kv_procedure.py:117-118: .float() of both inputs

    def train(self, B, steps, warmup):
        """(graphs/s, ms/step) of `steps` timed train steps after `warmup`"""
        V, A, y = self.batch(B)
        times = []
        for _ in range(warmup + steps):
            t0 = time.perf_counter()
            self._step(V, A, y)
            times.append(time.perf_counter() - t0)
        t = sum(times[warmup:]) / max(steps, 1)
        return B / t, t * 1e3

    def eval_forward(self, B, N, reps=1):
        """graphs/s of the eval forward at `N` nodes per graph (cfg5's CPU column)"""
        torch = self.torch
        from graph_representation_learning_b200 import synthetic
        wl = synthetic.Workload("cfg5", B, N)
        V, A, _ = synthetic.make_batch(wl, 77 + N, "cpu", B=B)
        V, A = V.contiguous(), A.float()
        if self.kind == "port":
            from oracle import reference_port as rp
            fwd = lambda: rp.forward(self._sd, V, A, training=False)  # noqa: E731
        else:
            self.model.eval()
            fwd = lambda: self.model([V, A])  # noqa: E731
        with torch.no_grad():
            fwd()
            t0 = time.perf_counter()
            for _ in range(reps):
                fwd()
        return B * reps / (time.perf_counter() - t0)


def run_reference(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    from graph_representation_learning_b200 import synthetic
    wl = synthetic.CONFIGS[args.workload]
    B = args.cpu_batch
    steps, warmup = max(1, args.steps), max(0, args.warmup)
    ref = CpuReference(wl)
    val, ms = ref.train(B, steps, warmup)
    sample = (f"{B} graphs x {wl.N} nodes per step ({warmup}+{steps} steps) of {wl.name}, dense fp32 adjacency, "
              f"{ref.cores} threads")
    extra = {}
    try:
        import psutil
        free_gb = psutil.virtual_memory().available / 1e9
    except Exception:
        free_gb = 0.0
    if wl.name.startswith("cfg2"):
        # a 1-thread run, and one step at the configuration's full batch when host RAM allows
        one = CpuReference(wl, threads=1)
        v1, ms1 = one.train(min(B, 4), 1, 1)
        extra["single_thread"] = {"value": v1, "unit": UNIT, "ms_per_step": ms1, "graphs_per_step": min(B, 4), "cores": 1}
        del one
        ref = CpuReference(wl)
        if free_gb > 64:
            vf, msf = ref.train(wl.B, 1, 0)
            extra["full_batch_step"] = {"value": vf, "unit": UNIT, "ms_per_step": msf, "graphs_per_step": wl.B,
                                        "cores": ref.cores, "note": "one un-warmed step at the configuration's batch"}
    print(json.dumps({
        "impl": "reference", "metric": METRIC, "value": val, "unit": UNIT, "n_gpus": args.gpus, "steps": steps,
        "warmup": warmup, "ms_per_step": ms, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
        "dtype": "f32", "data": "synthetic",
        "config": {"workload": wl.name, "step": "fwd+bwd+clip+adam", "graphs_per_step": B},
        "cpu_baseline": {"value": val, "unit": UNIT, "cores": ref.cores, "kind": ref.kind, "sample": sample},
        "e2e": {"value": val, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        **extra,
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
    if args.x3_promo:
        ops.set_x3_promotion(args.x3_promo)
    pk = peaks()

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def timed(fn, steps, warmup, sample_clocks=False):
        """(ms per step as the max over ranks, our kernel launches in the timed region, clocks)"""
        for _ in range(warmup):
            fn()
        barrier()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        n0 = _lib.launches()
        sampler = ClockSampler(local_rank) if (rank == 0 and sample_clocks) else None
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

    def make_step(wl):
        torch.manual_seed(0)                                     # identical replicas
        model = GraphCNNDropEdge(wl.input_dim, wl.output_dim, wl.num_edges, wl.net_size, wl.use_attention).to(dev)
        model.set_compute_dtype(wl.dtype)
        model.rng_seed = 1000 + rank                             # independent DropEdge/dropout streams per rank
        return model, KVTrainStep(model)

    def exposed_allreduce(step, fn, steps):
        """ms per step with the gradient collective skipped (None on one GPU): step time minus this = what the all-reduce
        costs on the critical path"""
        if world == 1:
            return None
        bucket = step.bucket
        bucket.all_reduce = lambda average=True: bucket.reattach()
        bucket.arm_overlap = lambda: None
        try:
            ms_off, _, _ = timed(fn, steps, 1)
        finally:
            del bucket.all_reduce, bucket.arm_overlap
        return ms_off

    wl = synthetic.CONFIGS[args.workload]
    B = args.batch or wl.B
    model, step = make_step(wl)
    grad_bytes = 4 * step.bucket.flat.numel()
    feat_dtype = torch.float32 if wl.dtype == torch.float32 else torch.bfloat16
    V, A, y = synthetic.make_batch(wl, 1234 + rank, dev, B=B, feature_dtype=feat_dtype)
    batch_dev = {"textline_encoding": V, "adjacency_matrix": A, "node_label": y}

    # (1) device-resident inputs: the whole forward([V, A]) API call incl. the CSR/CSC pack of the dense fp16 A
    ms_step, launches, clocks = timed(lambda: step.run_train_step(batch_dev), args.steps, args.warmup, sample_clocks=True)
    value = world * B / (ms_step * 1e-3)
    ms_nocomm = exposed_allreduce(step, lambda: step.run_train_step(batch_dev), max(5, args.steps // 2))

    if args.profile_step and not args.profile_workload:
        torch.cuda.synchronize()
        torch.cuda.profiler.start()
        step.run_train_step(batch_dev)
        torch.cuda.synchronize()
        torch.cuda.profiler.stop()

    # (2) end to end: pinned HOST buffers -> H2D inside the timed region, loss read back to the host.  Two host loaders of
    # the SAME batch: the EdgeListCollate format (features + (b, i, l, j) triples; the adjacency never exists densely,
    # SURVEY 8f N1) is the headline; the reference's dense fp16 [B, N, L, N] adjacency is reported beside it.
    e2e = None

    def pin(t):
        return torch.empty(t.shape, dtype=t.dtype, pin_memory=True).copy_(t)

    def nbytes(d, keys):
        return sum(d[k].numel() * d[k].element_size() for k in keys)

    pinned_loss = torch.zeros(64, dtype=torch.float32).pin_memory()

    def e2e_run(host, steps, warmup):
        """the public step API over a host loader: H2D of every step's inputs (pinned -> device, one batch ahead on
        a copy stream) and a D2H read of every step's loss, all inside the timed region"""
        from graph_representation_learning_b200.trainer import DevicePrefetcher
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

    def h2d_probe():
        """plain pinned -> device cudaMemcpyAsync of 1 GiB, all ranks at once: the box's H2D ceiling per rank"""
        src = torch.empty(1 << 30, dtype=torch.uint8, pin_memory=True)
        dst = torch.empty(1 << 30, dtype=torch.uint8, device=dev)
        dst.copy_(src, non_blocking=True)
        barrier()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(4):
            dst.copy_(src, non_blocking=True)
        e1.record()
        barrier()
        t = torch.tensor([e0.elapsed_time(e1)], device=dev)
        if world > 1:
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
        gbs = 4 * (1 << 30) / (float(t) * 1e-3) / 1e9
        del src, dst
        return gbs

    e_steps = max(3, min(args.steps, 10))
    if not args.no_e2e:
        Vp = V._base if V._base is not None else V          # the padded storage (16-byte row pitch)
        host_dense = {k: pin(v) for k, v in batch_dev.items()}
        Vp_h = pin(Vp)
        host_edges = {"textline_encoding": Vp_h[..., :wl.input_dim], "features_padded": Vp_h,
                      "node_label": host_dense["node_label"], "edges": pin(A.nonzero().to(torch.int32)),
                      "edge_weights": None, "graph_shape": (B, wl.N, wl.num_edges)}
        ms_edges = e2e_run(host_edges, e_steps, 2)
        ms_dense = e2e_run(host_dense, e_steps, 2)
        h2d_bytes = nbytes(host_edges, ["features_padded", "node_label", "edges"])
        probe = h2d_probe()
        e2e = {"value": world * B / (ms_edges * 1e-3), "unit": UNIT, "h2d_bytes_per_step": h2d_bytes,
               "d2h_bytes_per_step": 4, "ms_per_step": ms_edges, "steps": e_steps,
               "how": "KVTrainStep.run_train_step over DevicePrefetcher(host loader of EdgeListCollate batches): pinned "
                      "host V (fp32, all columns) / y / int32 (b,i,l,j) edge triples -> device one batch ahead on a copy "
                      "stream, CSR/CSC packed on the device, loss read back every step",
               "h2d_probe": {"per_rank_GBps": probe, "aggregate_GBps": probe * world,
                             "loader_per_rank_GBps": h2d_bytes / (ms_edges * 1e-3) / 1e9,
                             "how": "1 GiB pinned -> device cudaMemcpyAsync x4 on every rank at once (max over ranks): the "
                                    "box's concurrent H2D ceiling, beside what the loader moves per rank and step time"},
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
    roofline = None
    if not args.no_roofline:
        roofline = spmm_and_gemm_roofline(torch, ops, PackedGraphBatch, timed, pk, wl, B, A, dev)

    # (4) CPU baseline: the reference itself on a bounded sample of the same workload (rank 0, N=1 only)
    cpu = None
    if rank == 0 and world == 1 and not args.no_cpu_baseline:
        ref = CpuReference(wl)
        cb = args.cpu_batch
        v, ms_cpu = ref.train(cb, 3, 1)
        cpu = {"value": v, "unit": UNIT, "cores": ref.cores, "kind": ref.kind, "ms_per_step": ms_cpu,
               "sample": f"{cb} graphs x {wl.N} nodes per step (1+3 steps) of the same workload, dense fp32 adjacency"}
        del ref

    # (5) the other BASELINE configurations
    v_gb = V.numel() * V.element_size() / 1e9
    del step, model, batch_dev, V, A, y
    torch.cuda.empty_cache()
    extra = {}
    if not args.no_configs and args.workload == "cfg2" and args.batch is None:
        extra["cfg3"] = run_cfg3(torch, synthetic, make_step, timed, exposed_allreduce, world, rank, dev, args)
        torch.cuda.empty_cache()
        if world == 1:
            extra["cfg1"] = run_cfg1(torch, synthetic, make_step, timed, dev)
            torch.cuda.empty_cache()
            extra["cfg5"] = run_cfg5(torch, synthetic, GraphCNNDropEdge, PackedGraphBatch, dev,
                                     rank == 0 and not args.no_cpu_baseline)
            torch.cuda.empty_cache()
            extra["cfg4"] = run_cfg4(torch, synthetic, PackedGraphBatch, ops, make_step, timed, pk, dev, args)
            torch.cuda.empty_cache()

    if rank == 0:
        allreduce = None
        if world > 1:
            allreduce = {"bytes": grad_bytes, "ms_per_step_without": ms_nocomm, "exposed_ms": ms_step - ms_nocomm,
                         "how": "two buckets: everything but emb1 is reduced on NCCL's stream while emb1's weight-gradient "
                                "GEMM runs, then emb1; sum of gradients (the loss is normalised by the GLOBAL valid-node count)"}
        print(json.dumps({
            "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
            "ms_per_step": ms_step, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
            "dtype": "f32" if wl.dtype == torch.float32 else "bf16", "data": "synthetic",
            "config": {"workload": wl.name, "step": "fwd+bwd+clip+adam (forward includes the CSR pack of dense A and its typed transpose)",
                       "graphs_per_gpu": B, "global_batch": world * B, "parallelism": f"dp{world}",
                       "l2": "inputs (V %.2f GB) larger than L2" % v_gb, "features": "dense U[0,1)",
                       "host_cpus_rank0": (f"{len(numa_cpus)} GPU-local cores" if numa_cpus else "unbound")},
            "e2e": e2e, "gpu_launches": launches, "clocks": clocks, "roofline": roofline, "cpu_baseline": cpu,
            "allreduce": allreduce, "bow_features": bow, **extra,
        }))
    if world > 1:
        dist.destroy_process_group()


def spmm_and_gemm_roofline(torch, ops, PackedGraphBatch, timed, pk, wl, B, A, dev):
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
                "frac_nominal": achieved / NOMINAL_HBM_GBS, "peak_nominal": NOMINAL_HBM_GBS, "traffic": traffic,
                "peak_source": pk["source"], "ms_per_launch": ms_spmm, "algorithmic_bytes": alg_bytes}
    dYt = torch.randn(g.num_rows, S, device=dev, dtype=wl.dtype)
    dX = torch.empty(B * wl.N, S, device=dev, dtype=wl.dtype)
    ms_bwd, _, _ = timed(lambda: ops.spmm_bwd_(g, w, dYt, dX), 20, 5)
    alg_bwd = 4 * (g.num_nodes + 1) + 12 * g.nnz + dYt.numel() * es + dX.numel() * es
    # what the backward of the train step runs since r02 (ops._GraphConv): U = A_t^T dZ, the forward slab kernel over the
    # typed transpose of the batch with the layer's DropEdge weights; same algorithmic bytes as the forward aggregation
    gT, wT = g.typed_transpose(), g.transposed_weights(w)
    ms_agg_t, _, _ = timed(lambda: ops.spmm_fwd_(gT, wT, X, Y), 20, 5)
    roofline["bwd"] = {"kernel": "spmm_slab2_kernel over the typed transpose (U = A_t^T dZ; autograd of robust_gcn.py:45-47 "
                                 "as aggregate-then-project)",
                       "achieved": alg_bytes / (ms_agg_t * 1e-3) / 1e9,
                       "frac": alg_bytes / (ms_agg_t * 1e-3) / 1e9 / pk["hbm_gbs"],
                       "frac_nominal": alg_bytes / (ms_agg_t * 1e-3) / 1e9 / NOMINAL_HBM_GBS, "ms_per_launch": ms_agg_t,
                       "algorithmic_bytes": alg_bytes, "traffic": traffic}
    roofline["bwd_legacy"] = {"kernel": "spmm_rows_kernel<transposed> (project-then-aggregate; still serves layers with C > F)",
                              "achieved": alg_bwd / (ms_bwd * 1e-3) / 1e9,
                              "frac": alg_bwd / (ms_bwd * 1e-3) / 1e9 / pk["hbm_gbs"],
                              "frac_nominal": alg_bwd / (ms_bwd * 1e-3) / 1e9 / NOMINAL_HBM_GBS, "ms_per_launch": ms_bwd,
                              "algorithmic_bytes": alg_bwd, "traffic": traffic_bwd}
    del X, Y, dYt, dX, gT, wT
    # GEMMs are the larger share of the step (about 60 %): the gcn3 projection forward [Nt x S, K = 7 * 2S] and its dX,
    # timed alone.  Tensor-bound figure = EXECUTED tensor flops (fp32: three tf32 products forward, tf32 + double-K bf16
    # backward = 2 tf32-product equivalents; bf16: one product) against the dense peak of that operand type
    # (tf32 = half the measured bf16 figure).  "useful" = 2*M*N*K of the fp32-grade product.
    Kg = 14 * S
    Xg = torch.randn(B * wl.N, Kg, device=dev, dtype=wl.dtype)
    Wg = torch.randn(S, Kg, device=dev) / Kg ** 0.5
    dZg = torch.randn(B * wl.N, S, device=dev, dtype=wl.dtype)
    ms_gf, _, _ = timed(lambda: ops.gemm_nt([Xg], [Wg]), 10, 3)
    # dX of gcn3 since r02: per input, U [Nt, 7 S] against the regrouped weights [S, 7 S] (long-K; two such launches)
    Ug = torch.randn(B * wl.N, 7 * S, device=dev, dtype=wl.dtype)
    Wb = torch.randn(S, 7 * S, device=dev) / (7 * S) ** 0.5
    ms_gb, _, _ = timed(lambda: ops.gemm_nt([Ug], [Wb], fp32_mode=ops.GemmConfig.fp32_mode_bwd), 10, 3)
    useful = 2.0 * B * wl.N * S * Kg
    useful_b = 2.0 * B * wl.N * S * 7 * S
    fp32 = wl.dtype == torch.float32
    # these GEMMs are timed ALONE (a few ms of back-to-back launches): the burst figure of MEASURED_PEAKS.json applies, the
    # sustained one (a 4 s cuBLAS run at the power cap) is reported beside it
    peak_tf = pk["bf16_tflops"] / 2 if fp32 else pk["bf16_tflops"]
    peak_tf_sustained = pk["bf16_tflops_sustained"] / 2 if fp32 else pk["bf16_tflops_sustained"]
    prod_f, prod_b = (3.0, 2.0) if fp32 else (1.0, 1.0)
    roofline["gemm"] = {
        "bound": "tensor", "unit": "TFLOP/s", "peak": peak_tf,
        "peak_how": "measured burst cuBLAS bf16 (kernels timed alone)" + (" / 2 (tf32 operands)" if fp32 else ""),
        "peak_sustained": peak_tf_sustained,
        "fwd": {"kernel": "gemm_bc_2sm_kernel<x3> (gcn3 projection, three tf32 products, CTA pairs)" if fp32 else "gemm_nt_bf16_2sm_kernel (gcn3 projection, CTA pairs)",
                "shape_mnk": [B * wl.N, S, Kg], "ms_per_launch": ms_gf, "useful": useful / ms_gf / 1e9,
                "frac_useful": useful / ms_gf / 1e9 / peak_tf,
                "achieved": prod_f * useful / ms_gf / 1e9, "frac": prod_f * useful / ms_gf / 1e9 / peak_tf},
        "bwd_dx": {"kernel": "gemm_bc_2sm_kernel (gcn3 dX per input from U = A_t^T dZ, tf32 + bf16 correction product, CTA pairs)" if fp32 else "gemm_nt_bf16_2sm_kernel (gcn3 dX per input)",
                   "shape_mnk": [B * wl.N, S, 7 * S], "ms_per_launch": ms_gb, "useful": useful_b / ms_gb / 1e9,
                   "frac_useful": useful_b / ms_gb / 1e9 / peak_tf,
                   "achieved": prod_b * useful_b / ms_gb / 1e9, "frac": prod_b * useful_b / ms_gb / 1e9 / peak_tf},
        "note": "inside the step every GEMM runs at the board power cap (sw_power_cap, ~1000 W; DESIGN.md 4.2); frac counts "
                "executed tensor work (3 / 2 products per fp32-grade product), frac_useful the fp32-grade product itself"}
    return roofline


def run_cfg3(torch, synthetic, make_step, timed, exposed_allreduce, world, rank, dev, args):
    """BASELINE configs[2]: S=512, bf16, 1000-node graphs, data parallel.  Weak: 128 graphs per GPU.  Strong: global
    batch 1024 split over the ranks.  Each: ms/step (max over ranks), aggregate graphs/s, exposed all-reduce, peak memory."""
    wl = synthetic.CONFIGS["cfg3"]
    out = {"workload": "cfg3: S=512 bf16, 1000 nodes per graph, ~8 neighbours per edge type, dense features", "dtype": "bf16"}
    for mode, Bg in (("weak", wl.B), ("strong", 1024 // world)):
        torch.cuda.empty_cache()
        torch.cuda.reset_peak_memory_stats()
        model, step = make_step(wl)
        out["allreduce_bytes"] = 4 * step.bucket.flat.numel()
        V, A, y = synthetic.make_batch(wl, 4321 + rank, dev, B=Bg, feature_dtype=torch.bfloat16)
        batch = {"textline_encoding": V, "adjacency_matrix": A, "node_label": y}
        n = 5 if Bg <= 256 else 3
        ms, launches, _ = timed(lambda: step.run_train_step(batch), n, 2)
        ms_off = exposed_allreduce(step, lambda: step.run_train_step(batch), n)
        out[mode] = {"graphs_per_gpu": Bg, "global_batch": Bg * world, "ms_per_step": ms,
                     "graphs_per_s": world * Bg / (ms * 1e-3), "gpu_launches": launches,
                     "exposed_allreduce_ms": None if ms_off is None else ms - ms_off,
                     "peak_mem_GB": torch.cuda.max_memory_allocated() / 1e9, "scaling": mode}
        if args.profile_step and args.profile_workload == "cfg3" and mode == "weak":
            torch.cuda.synchronize()
            torch.cuda.profiler.start()
            step.run_train_step(batch)
            torch.cuda.synchronize()
            torch.cuda.profiler.stop()
        del model, step, V, A, y, batch
    return out


def run_cfg1(torch, synthetic, make_step, timed, dev):
    """BASELINE configs[0] on the GPU: 8 graphs x 200 nodes, in_dim 4369, S = 256, fp32 -- the reference's own CPU-runnable
    case, launch-bound here.  Eager KVTrainStep against trainer.GraphedTrainStep (forward + CE + backward, dense pack
    included, replayed as ONE CUDA graph; clip + Adam eager behind it)."""
    from graph_representation_learning_b200.trainer import GraphedTrainStep
    wl = synthetic.Workload("cfg1: 8 graphs x 200 nodes, ~8 nbrs/type, S=256, fp32", 8, 200)
    V, A, y = synthetic.make_batch(wl, 4321, dev)
    batch = {"textline_encoding": V, "adjacency_matrix": A, "node_label": y}
    out = {"workload": wl.name, "dtype": "f32"}
    for graphed in (False, True):
        model, step = make_step(wl)
        runner = GraphedTrainStep(step, wl.B, wl.N, wl.input_dim, wl.num_edges) if graphed else step
        ms, launches, _ = timed(lambda: runner.run_train_step(batch), 20, 5)
        key = "cuda_graph" if graphed else "eager"
        out[key] = {"ms_per_step": ms, "graphs_per_s": wl.B / (ms * 1e-3), "gpu_launches_per_step_outside_graph": launches / 20}
        del runner, step, model
        torch.cuda.empty_cache()
    return out


def run_cfg4(torch, synthetic, PackedGraphBatch, ops, make_step, timed, pk, dev, args):
    """BASELINE configs[3]: ONE graph, 1M nodes, 6 edge types, degree 16 per type, fp32, no attention: full-batch
    forward + backward.  The graph's feature slab does not fit in shared memory and the neighbours are uniform-random, so the
    gather kernels run and the SpMM roofline is `bytes_gather`: every entry fetches its own F*s-byte row."""
    wl = synthetic.CONFIGS["cfg4"]
    torch.cuda.reset_peak_memory_stats()
    model, step = make_step(wl)
    V, edges, y = synthetic.make_edge_batch(wl, 1234, dev)
    g = PackedGraphBatch.from_edges(edges, wl.B, wl.N, wl.num_edges)
    del edges
    batch = {"textline_encoding": V, "adjacency_matrix": g, "node_label": y}
    ms, launches, _ = timed(lambda: step.run_train_step(batch), 3, 1)
    if args.profile_step and args.profile_workload == "cfg4":
        torch.cuda.synchronize()
        torch.cuda.profiler.start()
        step.run_train_step(batch)
        torch.cuda.synchronize()
        torch.cuda.profiler.stop()
    out = {"workload": wl.name, "dtype": "f32", "ms_per_step": ms, "nodes_per_s": wl.N / (ms * 1e-3), "nnz": g.nnz,
           "gpu_launches": launches, "peak_mem_GB": torch.cuda.max_memory_allocated() / 1e9,
           "multi_gpu": "replicas only (one graph does not shard without a halo exchange)"}
    del model, step, V, batch
    torch.cuda.empty_cache()
    S = wl.net_size
    X = torch.randn(wl.N, S, device=dev)
    Y = torch.empty(g.num_rows, S, device=dev)
    w = g.edge_weights(0.3, seed=1, stream_id=1)
    ms_f, _, _ = timed(lambda: ops.spmm_fwd_(g, w, X, Y), 5, 2)
    # backward aggregation since r02: U = A_t^T dZ, the SAME kernel over the typed transpose of the graph (ops._GraphConv)
    gT, wT = g.typed_transpose(), g.transposed_weights(w)
    ms_b, _, _ = timed(lambda: ops.spmm_fwd_(gT, wT, X, Y), 5, 2)
    kept = int((w != 0).sum())
    idx = 4 * (g.num_rows + 1) + 8 * g.nnz
    gather_f = idx + kept * S * 4 + Y.numel() * 4                 # every KEPT entry fetches one F*4-byte row of X
    gather_b = gather_f
    tj = None
    tpath = os.path.join(ROOT, "profiles", "spmm_traffic_cfg4.json")
    if os.path.exists(tpath):
        tj = json.load(open(tpath))
    out["spmm"] = {
        "fwd": {"kernel": "spmm_nodes_kernel (L2 gather; the 1M x 1 KB slab does not fit shared memory)", "ms_per_launch": ms_f,
                "bytes_gather": gather_f, "achieved": gather_f / (ms_f * 1e-3) / 1e9, "unit": "GB/s",
                "frac": gather_f / (ms_f * 1e-3) / 1e9 / pk["hbm_gbs"], "frac_nominal": gather_f / (ms_f * 1e-3) / 1e9 / NOMINAL_HBM_GBS,
                "traffic": tj["fwd"]["dram_bytes_per_launch"] if tj else None},
        "bwd": {"kernel": "spmm_nodes_kernel over the typed transpose (U = A_t^T dZ, aggregate-then-project backward)", "ms_per_launch": ms_b, "bytes_gather": gather_b,
                "achieved": gather_b / (ms_b * 1e-3) / 1e9, "unit": "GB/s", "frac": gather_b / (ms_b * 1e-3) / 1e9 / pk["hbm_gbs"],
                "frac_nominal": gather_b / (ms_b * 1e-3) / 1e9 / NOMINAL_HBM_GBS,
                "traffic": tj["bwd"]["dram_bytes_per_launch"] if tj else None},
        "kept_entries": kept, "rule": "SURVEY 8d: cfg4 is judged on bytes_gather (no locality by construction); traffic = "
                                      "ncu dram bytes of the same launch (profiles/), ~ bytes_gather x (1 - L2 hit)"}
    return out


def run_cfg5(torch, synthetic, GraphCNNDropEdge, PackedGraphBatch, dev, cpu_column):
    """BASELINE configs[4]: eval forward at batch 64 over nodes per graph 100 -> 10k: latency p50 / p99 (one event pair per
    call, a synchronize between calls as a serving loop would see) and graphs/s, edge-list pack included."""
    wl0 = synthetic.Workload("cfg5", 64, 0)
    torch.manual_seed(0)
    model = GraphCNNDropEdge(wl0.input_dim, wl0.output_dim, wl0.num_edges, wl0.net_size, True).to(dev).eval()
    ref = CpuReference(wl0) if cpu_column else None
    rows = []
    for N in (100, 200, 500, 1000, 2000, 5000, 10000):
        torch.cuda.empty_cache()
        V, edges, _ = synthetic.make_edge_batch(wl0, 77 + N, dev, B=64, N=N)

        def fwd():
            with torch.no_grad():
                g = PackedGraphBatch.from_edges(edges, 64, N, wl0.num_edges, with_transpose=False)
                return model([V, g])
        for _ in range(2):
            fwd()
        lat = []
        for _ in range(10 if N >= 5000 else 30):
            torch.cuda.synchronize()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            fwd()
            e1.record()
            torch.cuda.synchronize()
            lat.append(e0.elapsed_time(e1))
        lat.sort()
        row = {"N": N, "latency_p50_ms": lat[len(lat) // 2], "latency_p99_ms": lat[min(len(lat) - 1, int(0.99 * len(lat)))],
               "graphs_per_s": 64 / (lat[len(lat) // 2] * 1e-3)}
        if ref is not None and N <= 2000:
            # the reference's dense CPU forward at a batch the host can hold, normalised to graphs/s
            Bc = 8 if N <= 500 else 2
            row["cpu_graphs_per_s"] = ref.eval_forward(Bc, N)
            row["cpu_batch"] = Bc
        rows.append(row)
        del V, edges
    return {"workload": "cfg5: eval forward, batch 64, nodes per graph 100 -> 10 000 (edge-list pack + model)", "dtype": "f32",
            "rows": rows, "cpu_kind": ref.kind if ref else None, "cpu_cores": ref.cores if ref else None}


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
