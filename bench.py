#!/usr/bin/env python
"""bench.py — megapixels/s of the Restormer forward (BASELINE.json configs[1]: bf16, 256x256, batch 16/GPU).

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl b200|reference]

One "step" = one forward of the default Restormer over one batch of 16 synthetic 256x256 RGB images per GPU
(batch-sharded, no data-path collective => weak scaling).  Prints ONE JSON line (rank 0).
  value     : whole-job MP/s, inputs already resident in HBM, CUDA-event timed, max over ranks.
  e2e       : same metric through the public nn.Module call with pinned-HOST input and output inside the
              timed region (H2D of the fp32 batch, forward, D2H of the restored fp32 batch).
  roofline  : dominant kernel (round 2: the fused LN + project_in + depthwise-3x3 + GELU gate at the full-resolution
              C=96 levels), algorithmic bytes / live CUDA-event duration vs MEASURED_PEAKS.json hbm_gbs.
  cpu_baseline / --impl reference: the oracle port (same ATen CPU ops as the reference's Restormer.py) on the
              host cores, bounded sample (B=1).
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

import torch  # noqa: E402

METRIC = "megapixels/sec (Restormer 256x256 fwd, bf16, batch 16 per GPU)"
H = W = 256
BATCH = 16
N_INPUT_SETS = 12  # 12 x 12.6 MB fp32 inputs (151 MB) > 126 MB L2; every level-1 activation is 100-268 MB anyway


def peaks():
    try:
        return json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json"))), "measured"
    except Exception:
        return {"hbm_gbs": 6650.0, "bf16_tflops": 1590.0, "bf16_tflops_sustained": 1400.0}, "fallback"


def ncu_traffic(kernel_tag="gate"):
    """dram read+write bytes per launch of the dominant kernel from the committed `ncu --set full` summary."""
    import glob
    unit = {"byte": 1.0, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9}
    for f in sorted(glob.glob(os.path.join(ROOT, "profiles", f"*_{kernel_tag}_ncu_full_summary.json")), reverse=True):
        try:
            d = json.load(open(f))
            tot = 0.0
            for k in ("dram__bytes_read.sum", "dram__bytes_write.sum"):
                v, u = d[k]
                tot += float(v) * unit[u]
            return tot, os.path.basename(f)
        except Exception:
            continue
    return None, None


class ClockSampler:
    """nvidia-smi clocks / throttle reasons DURING the timed region (B200_PROFILING.md recipe)."""
    Q = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
         "clocks_event_reasons.sw_power_cap")

    def __init__(self, index):
        self.index, self.rows, self.proc = index, [], None

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits",
                                          "-lms", "100", "-i", str(self.index)], stdout=subprocess.PIPE, text=True)
            threading.Thread(target=self._read, daemon=True).start()
        except Exception:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append([c.strip() for c in line.split(",")])

    def stop(self):
        if self.proc is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        self.proc.terminate()
        sm = sorted(int(r[1]) for r in self.rows if len(r) > 2 and r[1].isdigit())
        mx = [int(r[2]) for r in self.rows if len(r) > 2 and r[2].isdigit()]
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        reasons = set()
        for r in self.rows:
            for i, n in enumerate(names):
                if len(r) > 5 + i and r[5 + i].lower().startswith("active"):
                    reasons.add(n)
        return {"sm_mhz": sm[len(sm) // 2] if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "reasons": sorted(reasons), "samples": len(self.rows)}


def refuse_debug_env():
    """A number measured with an A/B library or a stage-dropping debug switch is not a bench number."""
    bad = sorted(k for k in os.environ if k.startswith("B200IR_") and k != "B200IR_COMPUTE_DTYPE")
    if bad:
        raise SystemExit(f"bench.py refuses to run with debug variables set: {bad}")


def cpu_reference_run(steps, warmup, sample_hw=256, x=None, keep_out=False):
    """The reference's CPU path (oracle port: identical ATen CPU ops) on all host cores; B=1 sample per step."""
    from oracle import restormer_oracle as O
    cores = os.cpu_count() or 1
    torch.set_num_threads(cores)
    sd = O.synth_state_dict({}, seed=0)
    if x is None:
        x = O.synth_image(1, sample_hw, sample_hw, seed=1234)
    y = None
    with torch.no_grad():
        for _ in range(warmup):
            y = O.restormer_forward(sd, x)
        t0 = time.perf_counter()
        for _ in range(steps):
            y = O.restormer_forward(sd, x)
        dt = (time.perf_counter() - t0) / max(steps, 1)
    r = ((sample_hw * sample_hw / 1e6) / dt, dt, cores, f"1 of the 16 images ({sample_hw}x{sample_hw}, fp32) per step")
    return r + (y,) if keep_out else r


def incumbent_gpu(O, sd, xs, steps):
    """Informational: what stock eager PyTorch gives on the SAME GPU for the same forward (the oracle's functional
    restatement of Restormer.py on CUDA tensors = the ATen/cuDNN/cuBLAS kernels the reference would launch): fp32 with
    TF32 off (the reference's precision) and under bf16 autocast.  Never the target, never in `value`."""
    out = {}
    sdd = {k: v.to(xs[0].device) for k, v in sd.items()}
    tf32 = (torch.backends.cudnn.allow_tf32, torch.backends.cuda.matmul.allow_tf32)
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    try:
        for name, auto in (("fp32_tf32_off", False), ("bf16_autocast", True)):
            torch.backends.cudnn.allow_tf32 = torch.backends.cuda.matmul.allow_tf32 = False
            try:
                with torch.no_grad(), torch.autocast("cuda", dtype=torch.bfloat16, enabled=auto):
                    O.restormer_forward(sdd, xs[0])
                    torch.cuda.synchronize()
                    e0.record()
                    for i in range(steps):
                        O.restormer_forward(sdd, xs[i % len(xs)])
                    e1.record()
                    torch.cuda.synchronize()
                ms = e0.elapsed_time(e1) / steps
                b, _, h, w = xs[0].shape
                out[name] = {"value": b * h * w / 1e6 / (ms * 1e-3), "unit": "MP/s", "ms_per_step": ms}
            except Exception as e:  # e.g. out of memory on a shared box: the leg is informational
                out[name] = {"error": str(e)[:200]}
                torch.cuda.empty_cache()
    finally:
        torch.backends.cudnn.allow_tf32, torch.backends.cuda.matmul.allow_tf32 = tf32
    out["what"] = "eager PyTorch (oracle restatement of Restormer.py on CUDA tensors), same batch, 1 GPU; informational"
    return out


def run_reference(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    steps, warmup = max(1, min(args.steps, 3)), max(1, min(args.warmup, 1))
    v, dt, cores, sample = cpu_reference_run(steps, warmup)
    print(json.dumps({
        "impl": "reference", "metric": METRIC, "value": v, "unit": "MP/s", "n_gpus": args.gpus, "steps": steps,
        "warmup": warmup, "ms_per_step": dt * 1e3, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
        "dtype": "f32", "data": "synthetic",
        "config": {"workload": "Restormer(dim=48,[4,6,6,8]) forward 256x256, CPU fp32 eager (oracle port of Restormer.py)"},
        "cpu_baseline": {"value": v, "unit": "MP/s", "cores": cores, "kind": "port", "sample": sample},
        "e2e": {"value": v, "unit": "MP/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
    }))


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--batch", type=int, default=BATCH)
    ap.add_argument("--dtype", default="bf16", choices=["bf16", "fp32"])
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-train", action="store_true", help="skip the cfg4 training-step leg")
    ap.add_argument("--no-hires", action="store_true", help="skip the cfg5 1024x1024 leg")
    ap.add_argument("--no-moce", action="store_true", help="skip the cfg3 MoCE-IR leg")
    ap.add_argument("--no-adair", action="store_true", help="skip the AdaIR leg")
    ap.add_argument("--kernel-table", default=None, help="write a per-kernel CUDA-event timing table (JSON) here")
    ap.add_argument("--no-incumbent", action="store_true", help="skip the eager-PyTorch-on-GPU informational leg")
    args = ap.parse_args()
    if args.impl == "reference":
        return run_reference(args)
    refuse_debug_env()
    args.warmup = max(args.warmup, 3)

    # NCCL's own log lines (e.g. 'NCCL version ...' when the box sets NCCL_DEBUG) go to stderr: stdout carries the JSON line only
    os.environ.setdefault("NCCL_DEBUG_FILE", "/dev/stderr")
    if os.environ.get("NCCL_DEBUG", "").upper() == "VERSION":   # (the VERSION level ignores NCCL_DEBUG_FILE)
        os.environ["NCCL_DEBUG"] = "WARN"
    import torch.distributed as dist
    import b200ir
    from b200ir import ops
    from oracle import restormer_oracle as O  # weights/inputs generator + cpu_baseline leg only

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    dev = torch.device("cuda", local)
    cd = torch.bfloat16 if args.dtype == "bf16" else torch.float32
    B = args.batch

    net = b200ir.Restormer().to(dev).eval().set_compute_dtype(cd)
    net.load_state_dict(O.synth_state_dict({}, seed=0), strict=True)
    host = [O.synth_image(B, H, W, seed=1234 + 100 * rank + i).pin_memory() for i in range(N_INPUT_SETS)]
    dev_in = [h.to(dev) for h in host]
    host_out = torch.empty((B, 3, H, W), dtype=torch.float32).pin_memory()

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def max_over_ranks(ms):
        if world == 1:
            return ms
        t = torch.tensor([ms], device=dev, dtype=torch.float64)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return t.item()

    mp_per_step_total = world * B * H * W / 1e6
    GATE_KEY = "dwconv3x3_gelu_gate"
    FUSED_KEY = "ln_proj_dw"
    with torch.no_grad():
        for i in range(args.warmup):
            net(dev_in[i % N_INPUT_SETS])
        # ---- device-resident timing -------------------------------------------------------
        sampler = ClockSampler(local)
        if rank == 0:
            sampler.start()
        barrier()
        l0 = ops.LAUNCHES
        ops.timing_start(only=[GATE_KEY, FUSED_KEY])
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for i in range(args.steps):
            net(dev_in[i % N_INPUT_SETS])
        e1.record()
        barrier()
        ms_eager = max_over_ranks(e0.elapsed_time(e1) / args.steps)
        gate_t = ops.timing_stop()
        # ---- device-resident, the serving call: the same forward replayed from ONE CUDA graph (b200ir.GraphedForward;
        # bitwise equal to the eager call, tests/test_restormer_gpu.py) -> `value`.  The eager loop above stays as
        # `eager_ms_per_step` and is where the per-kernel events of `roofline` are taken.
        fwd = b200ir.GraphedForward(net, dev_in[0])
        for i in range(args.warmup):
            fwd(dev_in[i % N_INPUT_SETS])
        barrier()
        l0 = ops.LAUNCHES
        e0.record()
        for i in range(args.steps):
            fwd(dev_in[i % N_INPUT_SETS])
        e1.record()
        barrier()
        ms = max_over_ranks(e0.elapsed_time(e1) / args.steps)
        launches = ops.LAUNCHES - l0
        del fwd
        # ---- end-to-end: pinned host in -> forward -> pinned host out ---------------------
        # (b200ir.PipelinedForward: H2D of batch i+1 and D2H of batch i-1 on copy streams overlap the graph replay of
        # batch i; every step's input comes from pinned host memory and every step's output lands in pinned host memory)
        pipe = b200ir.PipelinedForward(net, dev_in[0])
        host_outs = [torch.empty_like(host_out).pin_memory() for _ in range(2)]
        for i in range(3):
            pipe.submit(host[i % N_INPUT_SETS], host_outs[i % 2])
        pipe.synchronize()
        barrier()
        e0.record()
        for i in range(args.steps):
            pipe.submit(host[i % N_INPUT_SETS], host_outs[i % 2])
        pipe.join()
        e1.record()
        barrier()
        ms_e2e = max_over_ranks(e0.elapsed_time(e1) / args.steps)
        e2e_check = float((host_outs[(args.steps - 1) % 2] - net(dev_in[(args.steps - 1) % N_INPUT_SETS]).cpu()).abs().max())
        del pipe
        clocks = sampler.stop() if rank == 0 else None
        # ---- parity samples (checked against the CPU oracle below, outside every timed region) ----
        par_bf16 = net(dev_in[0])[0:1].float().cpu() if rank == 0 else None
        par_x128 = O.synth_image(1, 128, 128, seed=1234)
        par_fp32 = None
        if rank == 0:
            net.set_compute_dtype(torch.float32)
            par_fp32 = net(par_x128.to(dev)).cpu()      # BASELINE configs[0]: B1 128x128 fp32
            net.set_compute_dtype(cd)
        incumbent = None
        if rank == 0 and not args.no_incumbent:
            incumbent = incumbent_gpu(O, O.synth_state_dict({}, seed=0), dev_in[:3], 3)
        # ---- optional per-kernel table (separate pass, not part of any reported number) ----
        table = None
        if args.kernel_table and rank == 0:
            ops.timing_start()
            net(dev_in[0])
            torch.cuda.synchronize()
            table = ops.timing_stop()

    # ---- high-resolution leg (BASELINE configs[4]): 1024x1024 full frame, batch 1 per GPU ----------------------------
    hires = None
    if not args.no_hires:
        HR = 1024
        hin = [O.synth_image(1, HR, HR, seed=777 + 10 * rank + i).to(dev) for i in range(3)]
        with torch.no_grad():
            hfwd = b200ir.GraphedForward(net, hin[0])
            for i in range(args.warmup):
                hfwd(hin[i % 3])
            barrier()
            e0.record()
            for i in range(args.steps):
                hfwd(hin[i % 3])
            e1.record()
            barrier()
            del hfwd
        ms_hr = max_over_ranks(e0.elapsed_time(e1) / args.steps)
        hires = {"workload": f"Restormer forward, one {HR}x{HR} RGB frame per GPU, {args.dtype}, b200ir.GraphedForward; BASELINE configs[4]",
                 "value": world * HR * HR / 1e6 / (ms_hr * 1e-3), "unit": "MP/s", "ms_per_step": ms_hr}
        del hin

    # ---- MoCE-IR leg (BASELINE configs[2]): whole all-in-one model, 256x256 degraded inputs, batch 8, random-init experts --
    moce = None
    if not args.no_moce:
        from oracle.gen_golden_cases import MOCEIR_CFG3, synth_like
        MB = 8
        mnet = b200ir.MoCEIR(**MOCEIR_CFG3)
        mnet.load_state_dict(synth_like(mnet.state_dict(), 81), strict=True)
        mnet = mnet.to(dev).eval().set_compute_dtype(cd)
        min_ = [O.synth_image(MB, H, W, seed=555 + 10 * rank + i).to(dev) for i in range(4)]
        with torch.no_grad():
            for i in range(args.warmup):
                mnet(min_[i % 4])
            barrier()
            ml0 = ops.LAUNCHES
            e0.record()
            for i in range(args.steps):
                mnet(min_[i % 4])
            e1.record()
            barrier()
        ms_mo = max_over_ranks(e0.elapsed_time(e1) / args.steps)
        moce = {"workload": f"MoCEIR(dim=48,[4,6,6,8],dec[2,4,4],4 experts, top-1) forward, {H}x{W}, batch {MB}/GPU, {args.dtype}, "
                            f"eval; BASELINE configs[2]; the router's softmax/top-k on [B,4] logits and one host sync per "
                            f"decoder block are plain torch, everything tensor-sized is libb200ir",
                "value": world * MB * H * W / 1e6 / (ms_mo * 1e-3), "unit": "MP/s", "ms_per_step": ms_mo,
                "gpu_launches_per_step": (ops.LAUNCHES - ml0) // max(args.steps, 1)}
        if not args.no_train and world == 1:
            # the reference's own training objective for this model (MoCE-IR-main/src/train.py:59-72): L1 + 0.01 * balance
            # loss, AdamW, 128x128 patches, batch 8; eager (the dispatcher's host sync per decoder block cannot be captured)
            mnet.train()
            mopt = b200ir.AdamW(mnet.parameters(), lr=2e-4)
            mc = [torch.rand(MB, 3, 128, 128, generator=torch.Generator().manual_seed(900 + i)).to(dev) for i in range(2)]
            md = [(c + 25.0 / 255.0 * torch.randn_like(c)).clamp(0, 1) for c in mc]
            for i in range(max(args.warmup, 4)):     # (the first steps also learn the weight-pack table of the routed experts)
                b200ir.train_step(mnet, mopt, md[i % 2], mc[i % 2], balance_loss_weight=0.01)
            torch.cuda.synchronize()
            ml0 = ops.LAUNCHES
            nst = max(args.steps // 2, 4)
            e0.record()
            for i in range(nst):
                mloss = b200ir.train_step(mnet, mopt, md[i % 2], mc[i % 2], balance_loss_weight=0.01)
            e1.record()
            torch.cuda.synchronize()
            ms_mt = e0.elapsed_time(e1) / nst
            moce["train_step"] = {"workload": "MoCEIR training step (L1 + 0.01 * balance loss, AdamW), 128x128, batch 8, "
                                              f"{args.dtype}, eager, single GPU",
                                  "value": MB * 128 * 128 / 1e6 / (ms_mt * 1e-3), "unit": "MP/s", "ms_per_step": ms_mt,
                                  "gpu_launches_per_step": (ops.LAUNCHES - ml0) // nst, "loss": float(mloss)}
            del mopt, mc, md
        del mnet, min_
        torch.cuda.empty_cache()

    # ---- AdaIR leg (no BASELINE config): the Restormer U-Net + three FreModules, forward and training step ----
    adair = None
    if not args.no_adair and world == 1:
        from oracle.gen_golden_cases import synth_like
        AB = 8
        anet = b200ir.AdaIR()
        asd = synth_like(anet.state_dict(), 83)
        for k in asd:
            if k.endswith("para1"):                       # (init is zeros: the frequency branch would be multiplied away)
                asd[k] = torch.full_like(asd[k], 0.5)
        anet.load_state_dict(asd, strict=True)
        anet = anet.to(dev).eval().set_compute_dtype(cd)
        ain = [O.synth_image(AB, H, W, seed=655 + i).to(dev) for i in range(2)]
        afwd = b200ir.GraphedForward(anet, ain[0])
        for i in range(args.warmup):
            afwd(ain[i % 2])
        torch.cuda.synchronize()
        e0.record()
        for i in range(args.steps):
            afwd(ain[i % 2])
        e1.record()
        torch.cuda.synchronize()
        ms_ad = e0.elapsed_time(e1) / args.steps
        adair = {"workload": f"AdaIR(dim=48,[4,6,6,8]) forward, {H}x{W}, batch {AB}, {args.dtype}, b200ir.GraphedForward, single GPU",
                 "value": AB * H * W / 1e6 / (ms_ad * 1e-3), "unit": "MP/s", "ms_per_step": ms_ad,
                 "gpu_launches_per_step": afwd.launches_per_call}
        del afwd, ain
        if not args.no_train:
            anet.train()
            aopt = b200ir.AdamW(anet.parameters(), lr=2e-4)
            ac = [torch.rand(AB, 3, 128, 128, generator=torch.Generator().manual_seed(950 + i)).to(dev) for i in range(2)]
            adg = [(c + 25.0 / 255.0 * torch.randn_like(c)).clamp(0, 1) for c in ac]
            astep = b200ir.GraphedTrainStep(anet, aopt, adg[0], ac[0])
            for i in range(args.warmup):
                astep(adg[i % 2], ac[i % 2])
            torch.cuda.synchronize()
            e0.record()
            for i in range(args.steps):
                aloss = astep(adg[i % 2], ac[i % 2])
            e1.record()
            torch.cuda.synchronize()
            ms_at = e0.elapsed_time(e1) / args.steps
            adair["train_step"] = {"workload": f"AdaIR training step (L1, AdamW; AdaIR-main/train.py:38-46), 128x128, batch {AB}, "
                                               f"{args.dtype}, b200ir.GraphedTrainStep, single GPU",
                                   "value": AB * 128 * 128 / 1e6 / (ms_at * 1e-3), "unit": "MP/s", "ms_per_step": ms_at,
                                   "loss": float(aloss)}
            del astep, aopt, ac, adg
        del anet
        torch.cuda.empty_cache()

    # ---- training-step leg (BASELINE configs[3]): L1 + AdamW, 128x128 patches, batch 8 per GPU, DDP all-reduce --------
    train = None
    if not args.no_train:
        del dev_in
        torch.cuda.empty_cache()
        TB, TS = 8, 128
        tnet = b200ir.Restormer().to(dev).set_compute_dtype(cd)
        tnet.load_state_dict(O.synth_state_dict({}, seed=0), strict=True)
        opt = b200ir.AdamW(tnet.parameters(), lr=2e-4)
        tg = torch.Generator().manual_seed(4321 + rank)
        cleans = [torch.rand(TB, 3, TS, TS, generator=tg).to(dev) for _ in range(4)]
        degs = [(c + 25.0 / 255.0 * torch.randn(c.shape, generator=tg).to(dev)).clamp(0, 1) for c in cleans]
        step = b200ir.GraphedTrainStep(tnet, opt, degs[0], cleans[0])
        for i in range(args.warmup):
            step(degs[i % 4], cleans[i % 4])
        barrier()
        e0.record()
        for i in range(args.steps):
            loss = step(degs[i % 4], cleans[i % 4])
        e1.record()
        barrier()
        ms_tr = max_over_ranks(e0.elapsed_time(e1) / args.steps)
        train = {"workload": f"Restormer training step (L1 + AdamW lr=2e-4), {TS}x{TS} patches, batch {TB}/GPU, "
                             f"{'DDP sum all-reduce of the flat fp32 gradient buffer (NCCL)' if world > 1 else 'single GPU'}"
                             f"; BASELINE configs[3]; whole step replayed from one CUDA graph",
                 "value": world * TB * TS * TS / 1e6 / (ms_tr * 1e-3), "unit": "MP/s", "ms_per_step": ms_tr,
                 "gpu_launches_per_step": step.launches_per_step, "loss": float(loss.item()),
                 "allreduce_bytes_per_step": opt.numel * 4 if world > 1 else 0}

    if rank != 0:
        if world > 1:
            dist.destroy_process_group()
        return

    pk, pk_src = peaks()
    # dominant kernel = the timed kernel with the largest total time in the step.  Since round 2 that is the fused
    # LN + project_in + dw3x3 + GELU gate (K4+K5) at the full-resolution C=96 levels (8 of 44 blocks); shapes the fused
    # kernel does not take (C=192/384) still run the stand-alone gate kernel, which is timed too.
    key = max((k for k in gate_t), key=lambda k: gate_t[k][0] * gate_t[k][1])
    kms = gate_t[key][1]
    if key[0] == FUSED_KEY:
        _, kb, kc, khd, kh, kw, _g = key
        # SURVEY.md §8d, fused K4+K5: read the C-channel residual stream once, write the Hd gated channels once (bf16)
        alg_bytes = kb * kh * kw * (kc + khd) * 2
        kname, tag = f"ln_proj_dw_gate (LN+project_in+dw3x3+GELU gate) B{kb} C{kc} Hd{khd} {kh}x{kw}", "fused"
        # what the two kernels this one replaces had to move (K4: C + 2Hd, K5: 2Hd + Hd channels per pixel, bf16)
        pair_bytes = kb * kh * kw * ((kc + 2 * khd) + 3 * khd) * 2
    else:
        _, kb, khd, kh, kw, kdt = key
        esz = 2 if kdt == 1 else 4
        alg_bytes = kb * kh * kw * (2 * khd + khd) * esz  # read 2*Hd planes, write Hd planes (K5 = 2Hd+Hd)
        kname, tag = f"dwconv3x3_gelu_gate B{kb} Hd{khd} {kh}x{kw}", "gate"
    achieved = alg_bytes / (kms * 1e-3) / 1e9
    roofline = {"kernel": kname, "bound": "hbm", "achieved": achieved,
                "peak": pk["hbm_gbs"], "peak_source": pk_src, "unit": "GB/s", "frac": achieved / pk["hbm_gbs"],
                "traffic": ncu_traffic(tag)[0], "traffic_source": ncu_traffic(tag)[1], "launch_ms": kms,
                "launches_timed": gate_t[key][0], "alg_bytes_per_launch": alg_bytes,
                "share_of_step": gate_t[key][0] * kms / args.steps / ms_eager,
                **({"unfused_pair": {"alg_bytes": pair_bytes, "frac_at_this_time": pair_bytes / (kms * 1e-3) / 1e9 / pk["hbm_gbs"],
                                     "what": "bytes of the K4 + K5 pair this kernel replaces / this kernel's time / HBM peak: "
                                             "the pair would need this fraction of the roof to be as fast"}}
                   if key[0] == FUSED_KEY else {}),
                "timed_in": "the eager loop (eager_ms_per_step): one CUDA-event pair around every launch of the kernel"}
    out = {
        "metric": METRIC, "value": mp_per_step_total / (ms * 1e-3), "unit": "MP/s", "n_gpus": world,
        "steps": args.steps, "warmup": args.warmup, "ms_per_step": ms, "eager_ms_per_step": ms_eager,
        "higher_is_better": True, "scaling": "weak",
        "vs_baseline": None, "dtype": args.dtype, "data": "synthetic",
        "config": {"workload": f"Restormer(dim=48,[4,6,6,8],heads[1,2,4,8]) forward, {H}x{W} RGB, batch {B}/GPU, "
                               f"batch-sharded; BASELINE configs[1]",
                   "call": "b200ir.GraphedForward (the forward replayed from one CUDA graph, inputs resident in HBM); "
                           "eager_ms_per_step = the plain nn.Module call",
                   "l2": f"{N_INPUT_SETS} rotating input batches (151 MB) > 126 MB L2; activations 100-268 MB each",
                   "parallelism": f"dp{world} (no data-path collective)"},
        "e2e": {"value": mp_per_step_total / (ms_e2e * 1e-3), "unit": "MP/s", "ms_per_step": ms_e2e,
                "api": "b200ir.PipelinedForward(net, x).submit(host_in, host_out): pinned host -> device (copy stream), forward "
                       "(CUDA-graph replay), device -> pinned host (copy stream); copies of batches i+1 / i-1 overlap forward i",
                "max_abs_vs_eager": e2e_check,
                "h2d_bytes_per_step": B * 3 * H * W * 4, "d2h_bytes_per_step": B * 3 * H * W * 4},
        "gpu_launches": launches, "clocks": clocks, "roofline": roofline,
    }
    if train is not None:
        out["train_step"] = train
    if hires is not None:
        out["hires"] = hires
    if moce is not None:
        out["moce_ir"] = moce
    if adair is not None:
        out["adair"] = adair
    if incumbent is not None:
        out["incumbent_gpu"] = incumbent
    if not args.no_cpu_baseline:
        # the timed CPU sample IS image 0 of input batch 0: its output doubles as the parity reference for the GPU output
        v, dt, cores, sample, ref256 = cpu_reference_run(2, 1, x=host[0][0:1].clone(), keep_out=True)
        out["cpu_baseline"] = {"value": v, "unit": "MP/s", "cores": cores, "kind": "port", "sample": sample,
                               "s_per_step": dt}
        with torch.no_grad():
            sd0 = O.synth_state_dict({}, seed=0)
            ref128 = O.restormer_forward(sd0, par_x128)       # warm-up of this shape + the parity reference
            t0 = time.perf_counter()
            O.restormer_forward(sd0, par_x128)
            dt128 = time.perf_counter() - t0
        # BASELINE configs[0] exactly (B1 128x128 fp32, the reference's own CPU-runnable case), beside the cfg2 sample above
        out["cpu_baseline"]["cfg1"] = {"value": 128 * 128 / 1e6 / dt128, "unit": "MP/s", "s_per_step": dt128,
                                       "sample": "BASELINE configs[0]: one 128x128 image, fp32, whole forward"}
        out["parity"] = {
            "bf16_psnr_db": O.psnr(par_bf16, ref256), "bf16_config": f"cfg2 sample 0 of {B}, {H}x{W}, {args.dtype} GPU vs fp32 CPU oracle",
            "fp32_max_abs": float((par_fp32 - ref128).abs().max()), "fp32_config": "cfg1 B1 128x128, fp32 GPU path vs fp32 CPU oracle",
            "tolerance": {"bf16_psnr_db_min": 50.0, "fp32_max_abs_max": 1e-4}}
        out["parity"]["ok"] = bool(out["parity"]["bf16_psnr_db"] >= 50.0 and out["parity"]["fp32_max_abs"] <= 1e-4)
    if table is not None:
        rows = sorted(([list(k), n, m, n * m] for k, (n, m) in table.items()), key=lambda r: -r[3])
        json.dump({"total_ms": sum(r[3] for r in rows), "rows": rows}, open(args.kernel_table, "w"), indent=0)
    print(json.dumps(out))
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
