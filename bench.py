#!/usr/bin/env python
"""bench.py — particle-evals/sec of one full PSO step (G forward + fitness + pbest/gbest + update).

    python bench.py --gpus 1 --steps 20 --warmup 5            # this framework (libpsob.so, B200)
    python bench.py --impl reference --steps 3 --warmup 1     # the reference's CPU path (oracle port) on host cores
    torchrun ... bench.py --gpus N ...                        # one rank per GPU, particles sharded by rank

Workload at N=1: BASELINE.json configs[1] — DCGAN MNIST, Generator(100,1,64) + Discriminator(1,64),
random init (initialize_weights), swarm 65,536, discriminator-score fitness.  With N ranks every rank
owns 65,536 particles (weak scaling) and the ranks exchange one (4+D)-word record per iteration.
Prints ONE JSON line on rank 0.
"""
import argparse
import json
import os
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
for _p in (ROOT, os.path.join(ROOT, "gan-discovery-pso_b200")):
    if _p not in sys.path:
        sys.path.insert(0, _p)

# Rank 0 prints exactly ONE JSON line on stdout.  Libraries write there too (NCCL prints its version banner to
# stdout at NCCL_DEBUG >= VERSION), so file descriptor 1 is pointed at stderr for the whole run and the JSON line
# goes to the saved original descriptor.
os.environ.setdefault("NCCL_DEBUG_FILE", "/dev/stderr")
sys.stdout.flush()
_REAL_STDOUT = os.dup(1)
os.dup2(2, 1)


def emit(line: dict) -> None:
    os.write(_REAL_STDOUT, (json.dumps(line) + "\n").encode())

import numpy as np
import torch

METRIC = "particle-evals/sec (PSO step incl. G fwd + fitness)"
UNIT = "particle-evals/s"
D_LATENT = 100
PARTICLES_PER_GPU = 65536
WORKLOAD = "DCGAN MNIST swarm=65,536 per GPU, Generator(100,1,64)+Discriminator(1,64) random init, discriminator-score fitness"
# --workload assessor: the reference's OWN objective (configs[0]: ResNet-50 assessor, src/pso/util_cnn.py) at GPU scale
WORKLOAD_ASSESSOR = ("DCGAN MNIST swarm=65,536 per GPU, Generator(100,1,64) random init + ResNet-50 assessor (8 classes, seeded "
                     "trained-like parameters), the reference's Discovery.fitness (rescale, softmax pick, optimize_out_training)")


def peaks():
    path = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(path):
        with open(path) as f:
            p = json.load(f)
        return dict(hbm=p.get("hbm_gbs", 6650.0), tf_burst=p.get("bf16_tflops", 1590.0),
                    tf_sustained=p.get("bf16_tflops_sustained", 1400.0), source="measured")
    return dict(hbm=6650.0, tf_burst=1590.0, tf_sustained=1400.0, source="fallback")


def make_models(seed=42, oracle=False):
    """Synthetic weights of the BASELINE configs[1] networks with the reference's init (src/utils/util_dcgan.py:45-48).
    The GPU arm builds them with the product's own modules; only the CPU legs (cpu_baseline / --impl reference) use
    the oracle's torch modules (same constructors and draws, so both arms see the same weights)."""
    if oracle:
        from oracle import port as M
    else:
        from src.utils import util_dcgan as M
    torch.manual_seed(seed)
    gen = M.Generator(D_LATENT, 1, 64)
    disc = M.Discriminator(1, 64)
    M.initialize_weights(gen)
    M.initialize_weights(disc)
    return gen.eval(), disc.eval()


def seeded_fill(model, seed):
    """'Trained-like' synthetic ResNet parameters, independent of construction order: every state_dict entry drawn in
    sorted-key order from one seeded generator (conv N(0, 2/fan_in), BN gamma U(0.7, 1.3) / U(0.1, 0.4) for a block's last
    BN, beta and running_mean N(0, 0.1), running_var U(0.5, 1.5), fc N(0, 0.07)) -- same draws as the test fixtures."""
    gen = torch.Generator().manual_seed(seed)
    sd = model.state_dict()
    with torch.no_grad():
        for k in sorted(sd.keys()):
            v = sd[k]
            if k.endswith("num_batches_tracked"):
                continue
            if v.dim() == 4:
                v.copy_(torch.randn(v.shape, generator=gen) * (2.0 / (v.shape[1] * v.shape[2] * v.shape[3])) ** 0.5)
            elif k.endswith("running_var"):
                v.copy_(torch.rand(v.shape, generator=gen) + 0.5)
            elif k.endswith("running_mean"):
                v.copy_(torch.randn(v.shape, generator=gen) * 0.1)
            elif k.startswith("fc."):
                v.copy_(torch.randn(v.shape, generator=gen) * (0.07 if v.dim() == 2 else 0.1))
            elif k.endswith(".weight"):
                lo, hi = (0.1, 0.4) if ".bn3." in k else (0.7, 1.3)
                v.copy_(torch.rand(v.shape, generator=gen) * (hi - lo) + lo)
            else:
                v.copy_(torch.randn(v.shape, generator=gen) * 0.1)


def make_assessor(seed=11, oracle=False):
    """ResNet-50 assessor of configs[0] (8 classes, 1-channel 28x28) with synthetic parameters."""
    if oracle:
        from oracle import port
        m = port.Resnet([3, 4, 6, 3], 1, list(range(8)), 8).eval()
    else:
        from src.pso import util_cnn
        m = util_cnn.get_cnn("ResNet50", 1, iid_classes=list(range(8)), n_class=8).eval()
    seeded_fill(m, seed)
    return m


# ------------------------------------------------------------------------------------------------
# CPU baseline: the reference's per-particle loop (oracle port), bounded sample
# ------------------------------------------------------------------------------------------------

def cpu_reference_rate(n_particles=64, n_iters=2, seed=42, workload="dscore"):
    """Times the faithful restatement of Swarm.optimize (python loop over particles, batch-1 G and D / assessor,
    numpy fp64 state).  Returns (particle-evals/s, seconds, threads)."""
    from oracle import port
    gen, disc = make_models(seed, oracle=True)
    if workload == "assessor":
        obj = port.make_assessor_objective(gen, make_assessor(oracle=True), D_LATENT, class_index=0)
    else:
        obj = port.make_dscore_objective(gen, disc, D_LATENT)
    sw = port.PortSwarm(obj, num_particles=n_particles, n_iterations=n_iters, dim_space=D_LATENT, tolerance=1e-5,
                        w_inertia=0.73, w_cogn=1.496, w_soci=1.496, compute_diversity=False)
    np.random.seed(seed)
    t0 = time.perf_counter()
    sw.optimize(early_stopping=False)
    dt = time.perf_counter() - t0
    return n_particles * n_iters / dt, dt, torch.get_num_threads()


def run_reference(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    torch.set_num_threads(os.cpu_count() or 1)
    n_particles = 1024 if args.workload == "dscore" else 128
    for _ in range(args.warmup):
        cpu_reference_rate(32 if args.workload == "dscore" else 8, 1, workload=args.workload)
    times = []
    for _ in range(args.steps):
        rate, dt, threads = cpu_reference_rate(n_particles, 1, workload=args.workload)
        times.append(dt)
    ms = 1e3 * float(np.mean(times))
    value = n_particles / (ms / 1e3)
    nets_ = "G+D" if args.workload == "dscore" else "G + ResNet-50 assessor"
    sample = f"{n_particles} particles x 1 PSO iteration per step (per-particle batch-1 {nets_} forward, numpy fp64 update)"
    line = {
        "impl": "reference", "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": args.gpus, "steps": args.steps,
        "warmup": args.warmup, "ms_per_step": ms, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
        "dtype": "f32", "data": "synthetic",
        "config": {"workload": (WORKLOAD if args.workload == "dscore" else WORKLOAD_ASSESSOR).replace("swarm=65,536", f"swarm={args.particles:,}"), "sample": sample,
                   "host": "oracle/port.py PortSwarm (reference is pure Python; no GPU)"},
        "cpu_baseline": {"value": value, "unit": UNIT, "cores": torch.get_num_threads(), "kind": "port", "sample": sample},
        "e2e": {"value": value, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    emit(line)


# ------------------------------------------------------------------------------------------------
# clocks sampler
# ------------------------------------------------------------------------------------------------

class ClockSampler:
    Q = ("clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, index=0):
        self.index, self.samples, self.reasons, self.max_mhz = index, [], set(), None
        self._stop = threading.Event()
        self._t = None

    def _loop(self):
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        while not self._stop.is_set():
            try:
                out = subprocess.run(["nvidia-smi", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits", "-i", str(self.index)],
                                     capture_output=True, text=True, timeout=5).stdout.strip().split(",")
                self.samples.append(float(out[0]))
                self.max_mhz = float(out[1])
                for n, v in zip(names, out[2:]):
                    if v.strip().lower() == "active":
                        self.reasons.add(n)
            except Exception:
                pass
            self._stop.wait(0.1)

    def __enter__(self):
        self._t = threading.Thread(target=self._loop, daemon=True)
        self._t.start()
        return self

    def __exit__(self, *a):
        self._stop.set()
        self._t.join(timeout=6)

    def summary(self):
        return {"sm_mhz": float(np.median(self.samples)) if self.samples else None, "sm_max_mhz": self.max_mhz,
                "reasons": sorted(self.reasons), "samples": len(self.samples)}


# ------------------------------------------------------------------------------------------------
# GPU arm
# ------------------------------------------------------------------------------------------------

def run_psob(args):
    from psob import engine, dist as pdist, _ffi
    import torch.distributed as dist
    engine.require_cuda()
    rank, world, local = pdist.init_from_env("nccl")
    torch.cuda.set_device(local)
    dev = torch.device(f"cuda:{local}")
    P = args.particles
    first = rank * P
    gen, disc = make_models()
    gnet = engine.GeneratorNet(gen.state_dict(), P, device=dev)
    assessor = args.workload == "assessor"
    if assessor:
        anet = engine.AssessorNet(make_assessor().state_dict(), P, device=dev)
        fitness = engine.ResnetFitness(anet, 0, "optimize_out_training")
        dnet = None
    else:
        dnet = engine.DiscriminatorNet(disc.state_dict(), P, device=dev)
        fitness = engine.DScoreFitness(dnet)
    sw = engine.SwarmEngine(P, D_LATENT, generator=gnet, fitness=fitness, w_inertia=0.73, w_cogn=1.496,
                            w_soci=1.496, seed=42, index_offset=first, max_iterations=4096, device=dev, need_image_f32=False)
    g = torch.Generator(device=dev).manual_seed(42 + rank)
    sw.init_particles(generator=g)
    use_graph = (world == 1) and not args.no_graph

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    # ---- timed region: K whole iterations, state resident in HBM --------------------------------
    sw.run(args.warmup, use_graph=use_graph)
    barrier()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    with ClockSampler(local) as clocks:
        barrier()
        e0.record()
        sw.run(args.steps, use_graph=use_graph)
        e1.record()
        barrier()
    ms_total = torch.tensor([e0.elapsed_time(e1)], device=dev)
    if world > 1:
        dist.all_reduce(ms_total, op=dist.ReduceOp.MAX)
    ms_per_step = float(ms_total) / args.steps
    value = P * world / (ms_per_step / 1e3)

    # ---- per-kernel breakdown (eager, CUDA events on the launching stream) -----------------------
    stage_ms = {}

    def timed(name, fn, reps=5):
        evs = []
        for _ in range(reps):
            a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            a.record()
            fn()
            b.record()
            evs.append((a, b))
        torch.cuda.synchronize()
        stage_ms[name] = float(np.median([a.elapsed_time(b) for a, b in evs]))

    cur = sw.z
    for i, (layer, y, y32) in enumerate(zip(gnet.layers, gnet.acts, gnet.acts32)):
        if i == len(gnet.layers) - 1 and gnet.skip_f32 and y is not None:
            y32 = None                      # exactly what the timed step runs: D-score fitness reads the bf16 image only
        timed(layer.prog.name, lambda l=layer, c=cur, y=y, y32=y32: l.run(c, P, y, y32))
        cur = y
    if assessor:
        lib = _ffi.lib()
        timed("assessor_pack_rescale", lambda: anet.pack(gnet.image(), True))
        for op, layer in zip(anet.ops, anet.layers):
            src, dst = anet.buf[op.src], (anet.buf[op.dst] if op.dst else None)
            if op.kind == "maxpool":
                q = op.pool
                timed("a.maxpool", lambda s_=src, d_=dst, q=q: _ffi.check(lib.psob_maxpool_nhwc(
                    s_.data_ptr(), s_.stride(0), P, q["h"], q["w"], q["cp"], q["k"], q["stride"], q["pad"], d_.data_ptr(), d_.stride(0),
                    int(anet.tiled[op.dst]), engine._stream()), "maxpool"))
                continue
            res = anet.buf[op.res] if op.res else None
            timed(op.prog.name, lambda l=layer, s_=src, d_=dst, o=op, r=res: _ffi.check(lib.psob_layer_run_residual(
                l.handle, s_.data_ptr(), s_.stride(0), P, engine._ptr(d_), engine._ptr(anet.logits if o.dst32 else None),
                engine._ptr(r), r.stride(0) if r is not None else 0, anet.impl, engine._stream()), o.prog.name))
        timed("softmax_score_fitness", lambda: (engine.softmax_select(anet.logits, 0, fitness.prob),
                                                engine.score_fitness(fitness.prob, sw.fit, "optimize_out_training")))
        fit_progs_ = [op.prog for op in anet.ops if op.prog is not None]
        stage_ms["assessor_hbm_floor_ms"] = float(sum(p_.meta.get("hbm_bytes", 0) for p_ in fit_progs_) * P / (peaks()["hbm"] * 1e9) * 1e3)
        stage_ms["assessor_tensor_floor_ms"] = float(2.0 * anet.macs_exact * P / (peaks()["tf_sustained"] * 1e12) * 1e3)
        stage_ms["assessor_total"] = float(sum(v for k, v in stage_ms.items() if k.startswith("a.") or k == "assessor_pack_rescale"))
    else:
        for name, launch in dnet.launch_plan(gnet.image_padded_bf16()):      # fused head: d.c1, d.c2 (+ head dot), d.c3.finish
            timed(name, launch)
        timed("score_fitness", lambda: engine.score_fitness(dnet.acts32[-1], sw.fit, "optimize_out_training"))
    timed("pbest_gbest", sw.personal_and_global_best)
    timed("pso_update", sw.move)

    fit_programs = [op.prog for op in anet.ops if op.prog is not None] if assessor else dnet.programs
    progs = {p.name: p for p in gnet.programs + fit_programs}
    top = max((k for k in stage_ms if k in progs), key=lambda k: stage_ms[k])
    pk = peaks()
    flops = 2.0 * progs[top].macs_exact * P
    achieved_tf = flops / (stage_ms[top] * 1e-3) / 1e12
    traffic = None
    tpath = os.path.join(ROOT, "profiles", "ncu_traffic_latest.json")       # dram read+write bytes per launch, from `ncu --set full`
    if os.path.exists(tpath):
        with open(tpath) as f:
            traffic = json.load(f).get(top)
    roofline = {"kernel": f"pixel_gemm_umma_kernel[{top}]", "bound": "tensor", "achieved": achieved_tf, "peak": pk["tf_sustained"],
                "unit": "TFLOP/s", "frac": achieved_tf / pk["tf_sustained"], "traffic": traffic, "peak_source": pk["source"],
                "algorithmic_flops_per_launch": flops, "ms_per_launch": stage_ms[top]}
    hbm_bytes = float(progs[top].meta.get("hbm_bytes", 0)) * P
    if hbm_bytes / (pk["hbm"] * 1e9) > flops / (pk["tf_sustained"] * 1e12):
        # the dominant layer moves more bytes than it has math for (the assessor's 1x1 convs and shortcut adds):
        # its bound is HBM, algorithmic bytes = every input map once + the output once
        gbs = hbm_bytes / (stage_ms[top] * 1e-3) / 1e9
        roofline = {"kernel": f"pixel_gemm_umma_kernel[{top}]", "bound": "hbm", "achieved": gbs, "peak": pk["hbm"], "unit": "GB/s",
                    "frac": gbs / pk["hbm"], "traffic": traffic, "peak_source": pk["source"],
                    "algorithmic_bytes_per_launch": hbm_bytes, "algorithmic_flops_per_launch": flops, "ms_per_launch": stage_ms[top]}
    upd_bytes = 5.0 * D_LATENT * 4 * P
    upd_gbs = upd_bytes / (stage_ms["pso_update"] * 1e-3) / 1e9
    roofline_update = {"kernel": "pso_update_kernel", "bound": "hbm", "achieved": upd_gbs, "peak": pk["hbm"], "unit": "GB/s",
                       "frac": upd_gbs / pk["hbm"], "traffic": None, "algorithmic_bytes_per_launch": upd_bytes,
                       "ms_per_launch": stage_ms["pso_update"]}
    net_flops = 2.0 * (gnet.macs_exact + (anet if assessor else dnet).macs_exact) * P
    tensor_frac_step = net_flops / (ms_per_step * 1e-3) / 1e12 / pk["tf_sustained"]

    # ---- end to end through the public API with HOST buffers -------------------------------------
    # each step: H2D of the swarm's positions+velocities from pinned memory, one full iteration,
    # D2H of the fitness vector and the global best
    h_pos = torch.empty(P, D_LATENT, dtype=torch.float32).pin_memory()
    h_vel = torch.empty(P, D_LATENT, dtype=torch.float32).pin_memory()
    h_pos.copy_(sw.pos.cpu())
    h_vel.copy_(sw.vel.cpu())
    h_fit = torch.empty(P, dtype=torch.float32).pin_memory()
    h_best = torch.empty(D_LATENT + 1, dtype=torch.float32).pin_memory()

    # Two device staging buffers and a copy stream: the upload of step k+1's inputs and the read-back of step k's
    # results overlap step k's / k+1's kernels (every step still moves all of its bytes inside the timed region).
    cs = torch.cuda.Stream(device=dev)
    cs_out = torch.cuda.Stream(device=dev)       # read-backs on their own stream: they must not queue behind the next step's upload
    stage_pos = [torch.empty_like(sw.pos) for _ in range(2)]
    stage_vel = [torch.empty_like(sw.vel) for _ in range(2)]
    ev_h2d = [torch.cuda.Event() for _ in range(2)]
    ev_free = [torch.cuda.Event() for _ in range(2)]
    ev_done, ev_d2h = torch.cuda.Event(), torch.cuda.Event()

    def upload(k):
        with torch.cuda.stream(cs):
            cs.wait_event(ev_free[k % 2])
            stage_pos[k % 2][:, :D_LATENT].copy_(h_pos, non_blocking=True)
            stage_vel[k % 2][:, :D_LATENT].copy_(h_vel, non_blocking=True)
            ev_h2d[k % 2].record(cs)

    def e2e_run(n):
        main = torch.cuda.current_stream()
        for e in ev_free:
            e.record(main)
        ev_d2h.record(main)
        upload(0)
        for k in range(n):
            main.wait_event(ev_h2d[k % 2])
            sw.pos.copy_(stage_pos[k % 2])
            sw.vel.copy_(stage_vel[k % 2])
            ev_free[k % 2].record(main)
            if k + 1 < n:
                upload(k + 1)
            engine.pack_latent(sw.pos, sw.z)
            main.wait_event(ev_d2h)                 # the previous step's results have left sw.fit / gbest
            sw.step()
            ev_done.record(main)
            with torch.cuda.stream(cs_out):
                cs_out.wait_event(ev_done)
                h_fit.copy_(sw.fit, non_blocking=True)
                h_best[:D_LATENT].copy_(sw.gbest_pos[:D_LATENT], non_blocking=True)
                h_best[D_LATENT:].copy_(sw.gbest_val, non_blocking=True)
                ev_d2h.record(cs_out)
        main.wait_event(ev_d2h)

    for t in stage_pos + stage_vel:
        t.zero_()
    e2e_run(max(3, args.warmup))
    barrier()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    e2e_run(args.steps)
    b.record()
    barrier()
    e2e_ms = torch.tensor([a.elapsed_time(b)], device=dev)
    if world > 1:
        dist.all_reduce(e2e_ms, op=dist.ReduceOp.MAX)
    e2e_value = P * world / (float(e2e_ms) / args.steps / 1e3)

    # ---- CPU baseline (rank 0, N=1 only), bounded sample ------------------------------------------
    cpu = None
    if rank == 0 and world == 1 and not args.no_cpu_baseline:
        torch.set_num_threads(os.cpu_count() or 1)
        n_cpu = 256 if assessor else 2048
        cpu_reference_rate(8 if assessor else 32, 1, workload=args.workload)
        rate, dt, threads = cpu_reference_rate(n_cpu, 2, workload=args.workload)
        cpu = {"value": rate, "unit": UNIT, "cores": threads, "kind": "port",
               "sample": f"{n_cpu} particles x 2 iterations of the per-particle reference loop (oracle/port.py PortSwarm), {dt:.1f} s"}

    if rank == 0:
        # layers, score, pbest+candidate+apply, update (assessor: + pack, max pool, softmax pick)
        launches_per_step = len(gnet.layers) + (1 if gnet.programs[0].meta.get("ones_cols") else 0) + (len(anet.ops) + 2 if assessor else len(dnet.launch_plan(gnet.image_padded_bf16()))) + 1 + 3 + 1
        line = {
            "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
            "ms_per_step": ms_per_step, "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "bf16",
            "data": "synthetic",
            "config": {"workload": (WORKLOAD_ASSESSOR if assessor else WORKLOAD).replace("swarm=65,536", f"swarm={P:,}"), "particles_per_gpu": P, "latent_dim": D_LATENT, "parallelism": f"dp{world}",
                       "l2_policy": "inputs_exceed_l2 (per-step working set %s)" % ("~12 GB" if assessor else "~5 GB"), "cuda_graph": bool(use_graph),
                       "state_dtype": "f32", "rng": "philox4x32-10 in-kernel"},
            "roofline": roofline, "roofline_update": roofline_update,
            "tensor_frac_whole_step": tensor_frac_step,
            "stage_ms": stage_ms,
            "cpu_baseline": cpu,
            "e2e": {"value": e2e_value, "unit": UNIT, "h2d_bytes_per_step": 2 * P * D_LATENT * 4,
                    "d2h_bytes_per_step": P * 4 + (D_LATENT + 1) * 4},
            "gpu_launches": launches_per_step * args.steps,
            "clocks": clocks.summary(),
        }
        emit(line)
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--impl", default="psob", choices=["psob", "reference"])
    ap.add_argument("--particles", type=int, default=PARTICLES_PER_GPU, help="particles per GPU")
    ap.add_argument("--workload", default="dscore", choices=["dscore", "assessor"],
                    help="dscore: BASELINE configs[1] (headline); assessor: the reference's own ResNet-50 fitness at the same swarm size")
    ap.add_argument("--no-graph", action="store_true")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    args = ap.parse_args()
    if args.warmup < 3 and args.impl == "psob":
        args.warmup = 3
    if args.impl == "reference":
        run_reference(args)
    else:
        run_psob(args)


if __name__ == "__main__":
    main()
