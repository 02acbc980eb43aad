#!/usr/bin/env python
"""Benchmark of the forward hot path: images/sec of `net(x, inference=True)` for PIP-Net ConvNeXt-tiny (26x26 grid),
224^2, 768 prototypes, 200 classes, batch 64 per GPU (BASELINE.json configs[1]); weak scaling over N GPUs.

    python bench.py --gpus 1 --steps 20 --warmup 5                      # this package (sm_100a kernels)
    python bench.py --impl reference --gpus 1 --steps 20 --warmup 5     # reference arithmetic on the host cores
    python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 --master-port P \
        bench.py --gpus N --steps K --warmup W

Prints ONE JSON line (rank 0). See DESIGN.md "Measurement" for what each key means.
"""
from __future__ import annotations

import argparse
import json
import os
import statistics
import subprocess
import sys
import threading
import time

import torch

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

METRIC = "images/sec fwd (ConvNeXt-tiny PIP-Net, 224^2)"
UNIT = "images/s"
WORKLOAD = "PIP-Net ConvNeXt-tiny-26 224x224, 768 prototypes, 200 classes, batch 64 per GPU (BASELINE.json configs[1])"
FLOP_PER_IMAGE = 40.094e9  # 2*MAC of conv/linear layers (torch FlopCounterMode on the reference)
CLASSES, SIZE = 200, 224


def quiet():
    """The factories print configuration lines (like the reference); keep stdout for the JSON line."""
    import contextlib
    return contextlib.redirect_stdout(sys.stderr)


def load_peaks():
    path = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(path):
        with open(path) as f:
            d = json.load(f)
        return dict(hbm_gbs=d["hbm_gbs"], bf16_burst=d["bf16_tflops"], bf16_sustained=d["bf16_tflops_sustained"],
                    source="measured (MEASURED_PEAKS.json)")
    return dict(hbm_gbs=6650.0, bf16_burst=1590.0, bf16_sustained=1400.0, source="fallback (B200_PROFILING.md)")


def build_model(device, precision):
    from count_pipnet_b200 import synth
    from count_pipnet_b200.pipnet import pipnet as my_pipnet
    with quiet():
        model, _ = my_pipnet.get_pipnet(CLASSES, synth.make_args(net="convnext_tiny_26", num_features=0))
    synth.load_synth(model, seed=0)
    model = model.eval().to(device)
    model.cpn_precision = precision
    model.materialize_proto_map = False  # north star: the map is written only when the caller asks for it
    return model


def oracle_state_dict():
    """Seeded weights for the CPU arms, generated from key/shape only (no product module runs on that path)."""
    from count_pipnet_b200 import synth
    from count_pipnet_b200.pipnet import pipnet as my_pipnet
    with quiet():
        model, _ = my_pipnet.get_pipnet(CLASSES, synth.make_args(net="convnext_tiny_26", num_features=0))
    synth.load_synth(model, seed=0)
    return {k: v.detach() for k, v in model.state_dict().items()}


ORACLE_CFG = dict(model="pipnet", net="convnext_tiny_26")


def time_oracle(sd, images_per_step, steps, warmup):
    """Reference arithmetic (oracle port of pipnet/pipnet.py:31-41 + torchvision ConvNeXt) on the host cores."""
    import oracle
    from count_pipnet_b200 import synth
    x = synth.synth_images(images_per_step, SIZE, seed=1234)
    for _ in range(warmup):
        oracle.forward(sd, ORACLE_CFG, x, inference=True)
    t0 = time.perf_counter()
    for _ in range(steps):
        oracle.forward(sd, ORACLE_CFG, x, inference=True)
    dt = time.perf_counter() - t0
    return images_per_step * steps / dt, dt / steps * 1e3


class ClockSampler:
    """Samples nvidia-smi clocks / throttle reasons while the timed region runs."""
    Q = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, index):
        self.rows, self.proc, self.index = [], None, index
        self.t0 = self.t1 = None

    def __enter__(self):
        """Started BEFORE the warm-up: nvidia-smi needs 0.1-1 s to come up (longer on an 8-GPU box), the timed region
        of 20 steps lasts 85 ms. Rows are stamped on arrival; begin()/end() bracket the timed region."""
        if self.index is None:
            return self
        try:
            self.proc = subprocess.Popen(["nvidia-smi", f"--id={self.index}", f"--query-gpu={self.Q}",
                                          "--format=csv,noheader,nounits", "-lms", "10"],
                                         stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.thread = threading.Thread(target=self._read, daemon=True)
            self.thread.start()
        except Exception:
            self.proc = None
        return self

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append((time.perf_counter(), [c.strip() for c in line.split(",")]))

    def wait_first(self, timeout=5.0):
        t = time.perf_counter()
        while not self.rows and time.perf_counter() - t < timeout and self.proc is not None:
            time.sleep(0.01)

    def begin(self):
        self.t0 = time.perf_counter()

    def end(self):
        self.t1 = time.perf_counter()

    def __exit__(self, *a):
        if self.proc is not None:
            self.proc.terminate()
            try:
                self.proc.wait(timeout=2)
            except Exception:
                self.proc.kill()

    def summary(self):
        sm, mx, reasons = [], [], set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        rows = [r for t, r in self.rows if self.t0 is not None and self.t0 <= t <= self.t1 + 0.015]
        window = "timed region"
        if not rows and self.t0 is not None:
            # a very short timed region can fall between two samples: take the samples of the loaded period around it
            rows = [r for t, r in self.rows if self.t0 - 0.1 <= t <= self.t1 + 0.1]
            window = "timed region +- 0.1 s (device under the same load)"
        for r in rows:
            try:
                sm.append(float(r[0]))
                mx.append(float(r[1]))
                for n, v in zip(names, r[3:7]):
                    if v.lower().startswith("active"):
                        reasons.add(n)
            except Exception:
                continue
        if not sm:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": [], "samples": 0}
        return {"sm_mhz": statistics.median(sm), "sm_max_mhz": max(mx), "reasons": sorted(reasons), "samples": len(sm),
                "window": window}


def base_line(args, world):
    return {"metric": METRIC, "unit": UNIT, "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
            "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "data": "synthetic",
            "config": {"workload": WORKLOAD, "batch_per_gpu": args.batch, "global_batch": args.batch * world,
                       "image_size": SIZE, "prototypes": 768, "classes": CLASSES,
                       "weights": "seeded synthetic (count_pipnet_b200/synth.py), random init shape of ConvNeXt-tiny"}}


# ------------------------------------------------------------------------------------------------ reference arm
def run_reference(args, rank, world):
    """The reference has no installable package and is pure PyTorch; its CPU implementation of this path cannot travel
    to the GPU box, so this arm times the oracle port of it (oracle/, pinned bit-exactly to the real reference by
    tests/golden) on the host cores with all threads. Rank 0 only."""
    if rank != 0:
        return
    torch.set_num_threads(os.cpu_count() or 1)
    sd = oracle_state_dict()
    rate, _ = time_oracle(sd, 2, 1, 1)  # probe, then size a bounded per-step sample: whole run about two minutes
    per_step = max(1, min(args.batch, int(rate * 120.0 / max(args.steps + args.warmup, 1))))
    value, ms = time_oracle(sd, per_step, args.steps, args.warmup)
    line = base_line(args, world)
    line.update({"impl": "reference", "value": value, "ms_per_step": ms, "dtype": "f32",
                 "cpu_baseline": {"value": value, "unit": UNIT, "cores": torch.get_num_threads(), "kind": "port",
                                  "sample": f"{per_step} images per step x {args.steps} steps of the same workload, "
                                            f"oracle/ (CPU restatement of the reference forward), torch {torch.__version__}"},
                 "e2e": {"value": value, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}})
    line["config"]["reference_sample_images_per_step"] = per_step
    print(json.dumps(line), file=JSON_OUT, flush=True)


# ------------------------------------------------------------------------------------------------ this package
def run_ours(args, rank, world, local_rank):
    import torch.distributed as dist
    from count_pipnet_b200 import engine
    device = torch.device("cuda", local_rank)
    torch.cuda.set_device(device)
    if world > 1:
        dist.init_process_group("nccl", device_id=device)
    model = build_model(device, args.precision)
    from count_pipnet_b200 import synth
    B = args.batch
    nrot = 8  # 8 resident batches = 308 MB at B=64 > 126 MB L2: inputs never come from cache
    xs = [synth.synth_images(B, SIZE, seed=1000 + rank * 100 + i).to(device) for i in range(nrot)]
    from count_pipnet_b200 import sharding

    def step(i):
        with torch.no_grad():
            _, pooled, out = model(xs[i % nrot], inference=True)
        if world > 1:  # the only exchange of the path: gather logits and pooled scores
            pooled, out = sharding.gather_outputs(pooled, out, B * world)
        return pooled, out

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    with ClockSampler(local_rank if rank == 0 else None) as clocks:  # rank 0 reports its own GPU
        for i in range(args.warmup):
            step(i)
        clocks.wait_first()
        for i in range(2):  # back under load after the wait (untimed)
            step(i)
        barrier()
        clocks.begin()
        ev0.record()
        for i in range(args.steps):
            step(i)
        ev1.record()
        barrier()
        clocks.end()
    ms_total = torch.tensor([ev0.elapsed_time(ev1)], device=device)
    if world > 1:
        dist.all_reduce(ms_total, op=dist.ReduceOp.MAX)
    ms_step = float(ms_total.item()) / args.steps
    value = world * B / (ms_step / 1e3)

    # ---- end to end: pinned host batches -> H2D -> forward -> D2H of logits + pooled, copy overlapped on a side stream
    host = [synth.synth_images(B, SIZE, seed=2000 + i).pin_memory() for i in range(4)]
    dev_in = [torch.empty(B, 3, SIZE, SIZE, device=device) for _ in range(2)]
    host_logits = torch.empty(B, CLASSES).pin_memory()
    host_pooled = torch.empty(B, 768).pin_memory()
    copy_stream = torch.cuda.Stream(device)
    main_stream = torch.cuda.current_stream(device)

    def e2e_loop(n):
        ready = [torch.cuda.Event() for _ in range(2)]
        consumed = [torch.cuda.Event() for _ in range(2)]
        for e in consumed:
            e.record(main_stream)
        for i in range(n):
            s = i % 2
            with torch.cuda.stream(copy_stream):
                copy_stream.wait_event(consumed[s])
                dev_in[s].copy_(host[i % len(host)], non_blocking=True)
                ready[s].record(copy_stream)
            main_stream.wait_event(ready[s])
            with torch.no_grad():
                _, pooled, out = model(dev_in[s], inference=True)
            consumed[s].record(main_stream)
            host_logits.copy_(out, non_blocking=True)
            host_pooled.copy_(pooled, non_blocking=True)
        torch.cuda.synchronize()

    e2e_loop(max(2, args.warmup))
    barrier()
    t0 = time.perf_counter()
    e2e_loop(args.steps)
    e2e_s = torch.tensor([time.perf_counter() - t0], device=device)
    if world > 1:
        dist.all_reduce(e2e_s, op=dist.ReduceOp.MAX)
    e2e_value = world * B * args.steps / float(e2e_s.item())

    line = base_line(args, world)
    line.update({"value": value, "ms_per_step": ms_step, "dtype": "bf16" if args.precision == "bf16" else "bf16x3",
                 "clocks": clocks.summary(),
                 "e2e": {"value": e2e_value, "unit": UNIT, "h2d_bytes_per_step": B * 3 * SIZE * SIZE * 4,
                         "d2h_bytes_per_step": B * (CLASSES + 768) * 4}})
    line["config"].update({"precision_mode": args.precision, "parallelism": f"dp{world} (batch sharded, no data-path "
                           "collective except the logits/pooled all-gather)",
                           "l2_policy": f"inputs rotate over {nrot} resident batches ({nrot * B * 3 * SIZE * SIZE * 4 >> 20} MiB > L2); "
                                        "activations per step exceed L2",
                           "proto_map": "not materialised (caller did not ask for it)"})

    if rank == 0:
        st = engine.state_for(model, device)
        rec = next(iter(st.plans.values()))
        line["gpu_launches"] = len(rec.plan) * args.steps
        line.update(roofline_report(rec.plan, B))
        if world == 1:
            torch.set_num_threads(os.cpu_count() or 1)
            sd = oracle_state_dict()
            cpu_val, _ = time_oracle(sd, 16, 4, 1)
            line["cpu_baseline"] = {"value": cpu_val, "unit": UNIT, "cores": torch.get_num_threads(), "kind": "port",
                                    "sample": "16 images x 4 timed forwards (1 warm-up) of the same workload through "
                                              "oracle/ (CPU restatement of the reference forward)"}
        print(json.dumps(line), file=JSON_OUT, flush=True)
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()


# dram__bytes_read.sum + dram__bytes_write.sum of one launch from an `ncu --set full` capture (not measured live)
NCU_TRAFFIC = {
    "mlp_fused": {"bytes": 137.6e6, "note": "stage-3 launch (M=46656, C=384), profiles/r01_mlp_fused.md: 114.5 MB read "
                  "+ 23.1 MB written by kernel end; algorithmic 112 MB in + 72 MB out (output still dirty in L2)"},
}


def roofline_report(plan, batch):
    """Per-kernel device times from an event-separated replay of the recorded step (3 passes, averaged), and the
    roofline of the dominant kernel family (the tcgen05 GEMM) from its algorithmic FLOPs."""
    peaks = load_peaks()
    agg = {}
    passes = 3
    for _ in range(passes):
        for name, ms, flops, nbytes in plan.run_timed():
            a = agg.setdefault(name, [0.0, 0.0, 0.0, 0])
            a[0] += ms; a[1] += flops; a[2] += nbytes; a[3] += 1
    total_ms = sum(a[0] for a in agg.values()) / passes
    kernels = {}
    for name, (ms, flops, nbytes, n) in agg.items():
        kernels[name] = {"launches_per_step": n // passes, "ms_per_step": ms / passes, "share": ms / passes / total_ms,
                         "tflops": flops / ms / 1e9 if ms > 0 else None, "gbs": nbytes / ms / 1e6 if ms > 0 else None}
    dom = max(kernels, key=lambda k: kernels[k]["ms_per_step"])
    is_tensor = lambda k: k.startswith("gemm") or k.startswith("mlp_fused")  # tcgen05 kernels
    if kernels[dom]["tflops"] and is_tensor(dom):
        roof = {"kernel": dom, "bound": "tensor", "achieved": kernels[dom]["tflops"], "peak": peaks["bf16_sustained"],
                "unit": "TFLOP/s", "frac": kernels[dom]["tflops"] / peaks["bf16_sustained"],
                "traffic": NCU_TRAFFIC.get(dom, {}).get("bytes"), "traffic_note": NCU_TRAFFIC.get(dom, {}).get("note"),
                "peak_source": peaks["source"] + ", sustained figure (kernel timed inside the step)"}
    else:
        roof = {"kernel": dom, "bound": "hbm", "achieved": kernels[dom]["gbs"], "peak": peaks["hbm_gbs"], "unit": "GB/s",
                "frac": kernels[dom]["gbs"] / peaks["hbm_gbs"], "traffic": None, "peak_source": peaks["source"]}
    hbm, tensor = {}, {}
    dwk = "dwconv7_stats" if "dwconv7_stats" in kernels else "dwconv7_ln"  # bf16 mode / fp32 mode kernel
    if dwk in kernels:
        hbm = {"kernel": dwk, "bound": "hbm", "achieved": kernels[dwk]["gbs"],
               "peak": peaks["hbm_gbs"], "unit": "GB/s", "frac": kernels[dwk]["gbs"] / peaks["hbm_gbs"],
               "note": "HBM by algorithmic bytes (4 B in + 2 B out per element); 49 fp32 FMAs per element make the "
                       "FMA pipe the practical bound (tflops in `kernels`, 74 TFLOP/s fp32 peak)"}
    fam = [k for k in agg if is_tensor(k)]
    if fam:
        # all tcgen05 kernels of the step together (fused CNBlock MLPs + the remaining GEMMs): FLOPs / device time
        f_ms = sum(agg[k][0] for k in fam)
        f_tf = sum(agg[k][1] for k in fam) / f_ms / 1e9
        tensor = {"kernel": "+".join(sorted(fam)), "bound": "tensor", "achieved": f_tf, "peak": peaks["bf16_sustained"],
                  "unit": "TFLOP/s", "frac": f_tf / peaks["bf16_sustained"], "ms_per_step": f_ms / passes}
    return {"roofline": roof, "roofline_hbm_kernel": hbm, "roofline_tensor_kernel": tensor, "kernels": kernels,
            "step_tflops": FLOP_PER_IMAGE * batch / (total_ms * 1e-3) / 1e12}


JSON_OUT = sys.stdout


def main():
    # stdout carries exactly ONE line, the JSON: libraries that print there (NCCL's version banner under torchrun, the
    # reference-style "Number of prototypes" message of the factories) are sent to stderr for the length of the run
    global JSON_OUT
    sys.stdout.flush()
    JSON_OUT = os.fdopen(os.dup(1), "w")
    os.dup2(2, 1)
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--batch", type=int, default=64, help="images per GPU per step")
    ap.add_argument("--precision", default="bf16", choices=["bf16", "fp32"])
    args = ap.parse_args()
    args.warmup = max(args.warmup, 3) if args.impl == "ours" else args.warmup
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    if args.impl == "reference":
        run_reference(args, rank, world)
    else:
        run_ours(args, rank, world, local_rank)


if __name__ == "__main__":
    main()
