#!/usr/bin/env python
"""bench.py — particle-evals/sec of one full PSO step (G forward + fitness + pbest/gbest + update).

    python bench.py --gpus 1 --steps 20 --warmup 5            # this framework (libpsob.so, B200)
    python bench.py --impl reference --steps 3 --warmup 1     # CPU arm on the host cores (batched restatement of this workload
                                                              #   + the unmodified reference on the workload it can run)
    torchrun ... bench.py --gpus N ...                        # one rank per GPU, particles sharded by rank (weak scaling)
    torchrun ... bench.py --gpus N --total-particles 1048576  # BASELINE configs[4]: ONE swarm of 1 M sharded (strong scaling)
    python bench.py --workload assessor|config3|config4       # the other BASELINE configs as bench lines

Workload at N=1: BASELINE.json configs[1] — DCGAN MNIST, Generator(100,1,64) + Discriminator(1,64),
random init (initialize_weights), swarm 65,536, discriminator-score fitness.  With N ranks every rank
owns 65,536 particles (weak scaling) and the ranks exchange one (4+D)-word record per iteration; before the timed
region a sharded-vs-unsharded trajectory check runs on the NCCL ranks themselves ("shard_parity").
Prints ONE JSON line on rank 0; exits 3 if the shard parity check fails.
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
WORKLOADS = {
    # BASELINE configs[1] (headline) and, with --total-particles 1048576 under torchrun, configs[4]
    "dscore": "DCGAN MNIST swarm=65,536 per GPU, Generator(100,1,64)+Discriminator(1,64) random init, discriminator-score fitness",
    # the reference's OWN objective (configs[0]: ResNet-50 assessor, src/pso/util_cnn.py) at GPU scale
    "assessor": ("DCGAN MNIST swarm=65,536 per GPU, Generator(100,1,64) random init + ResNet-50 assessor (8 classes, seeded "
                 "trained-like parameters), the reference's Discovery.fitness (rescale, softmax pick, optimize_out_training)"),
    # BASELINE configs[2] / configs[3]: parity-test cases that can also print a line
    "config3": "VQ-VAE-GAN decoder (VectorQuantizedVAE_GAN.decoder, xavier init) swarm=16,384 per GPU, per-particle fp32 28x28 targets, reconstruction-loss fitness",
    "config4": "G256(z=100,f=64) (extension) swarm=8,192 per GPU, per-particle bf16 1x256x256 targets, inverter-loss fitness (w_ass=0)",
}
DEFAULT_PARTICLES = {"dscore": 65536, "assessor": 65536, "config3": 16384, "config4": 8192}


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
# CPU legs (the only code in this file that may execute oracle/):
#   reference    the UNMODIFIED reference classes (oracle/_ref, copied by oracle/make_ref.sh) -- BASELINE config 1, the only
#                workload the reference itself can run (ResNet-50 assessor; a Discriminator as model_cnn raises IndexError)
#   restatement  batched fp32 torch-CPU restatement of the D-score iteration (oracle/port.py BatchedDScoreSwarm) -- the
#                like-for-like CPU figure for configs 2 / 5; never labelled "reference"
# ------------------------------------------------------------------------------------------------

def cpu_reference_config1(n_iterations=10):
    """-> cpu_baseline dict with kind "reference", or None when oracle/_ref is absent."""
    from oracle import run_reference
    if run_reference.reference_root() is None:
        return None
    r = run_reference.time_config1(num_particles=64, n_iterations=n_iterations)
    return {"value": r["evals_per_s"], "unit": UNIT, "cores": r["threads"], "kind": "reference",
            "sample": f"unmodified reference Swarm.optimize ({r['root']}/src/pso/util_pso.py:167-193), BASELINE config 1: Generator(100,1,64) + "
                      f"ResNet-50 assessor, P=64 x {n_iterations} iterations, early_stopping=False, {r['seconds']:.1f} s "
                      "(the only workload the reference can run: a Discriminator as model_cnn raises IndexError at util_discovery.py:74)",
            "breakdown_s": r["breakdown_s"], "cpu_count": r["cpu_count"], "torch": r["torch"], "numpy": r["numpy"]}


def cpu_restatement_dscore(n_particles=4096, n_iters=2, seed=42, chunk=2048):
    """Batched fp32 torch-CPU restatement of the bench's own workload.  Returns (evals/s, seconds, threads)."""
    from oracle import port
    gen, disc = make_models(seed, oracle=True)
    sw = port.BatchedDScoreSwarm(gen, disc, n_particles, D_LATENT, 0.73, 1.496, 1.496, seed=seed, chunk=chunk)
    sw.step()                                   # warm-up (thread pools, oneDNN primitive caches)
    t0 = time.perf_counter()
    for _ in range(n_iters):
        sw.step()
    dt = time.perf_counter() - t0
    return n_particles * n_iters / dt, dt, torch.get_num_threads()


def cpu_port_rate(n_particles=64, n_iters=2, seed=42, workload="assessor"):
    """Per-particle restatement of Swarm.optimize (oracle/port.py PortSwarm) -- fallback when oracle/_ref is absent."""
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
    """`--impl reference`: the CPU arm on this box's host cores, rank 0 only.
    dscore (default; the GPU arm's own config): the batched restatement, honestly labelled -- plus, beside it, the unmodified
    reference on the one workload it can run.  assessor: the unmodified reference itself (config 1)."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    torch.set_num_threads(os.cpu_count() or 1)
    if args.workload not in ("dscore", "assessor"):
        emit({"impl": "reference", "unavailable": f"the reference has no implementation of workload {args.workload} (extension configs)"})
        return
    times, extra = [], {}
    if args.workload == "dscore":
        n_particles, iters = 4096, 1
        for _ in range(max(1, args.warmup)):
            cpu_restatement_dscore(512, 1)
        for _ in range(args.steps):
            rate, dt, threads = cpu_restatement_dscore(n_particles, iters)
            times.append(dt / iters)
        sample = (f"{n_particles} particles x {iters} PSO iteration per step: batched fp32 torch-CPU RESTATEMENT (oracle/port.py "
                  "BatchedDScoreSwarm: batched G+D forward, vectorised pbest / argmin / Philox move); the reference itself cannot "
                  "run a Discriminator as model_cnn (IndexError at src/pso/util_discovery.py:74)")
        kind, label = "port", "restatement"
        try:
            extra["reference_config1"] = cpu_reference_config1(n_iterations=5)
        except Exception as e:      # the reference copy is optional on this arm
            extra["reference_config1"] = {"unavailable": repr(e)[:200]}
    else:
        from oracle import run_reference as rr
        n_particles = 64
        if rr.reference_root() is not None:
            for _ in range(args.steps):
                r = rr.time_config1(num_particles=n_particles, n_iterations=2)
                times.append(r["seconds"] / 2)
            sample = f"unmodified reference Swarm.optimize ({r['root']}), BASELINE config 1, P=64 x 2 iterations per step"
            kind, label = "reference", "reference"
            extra["breakdown_s"] = r["breakdown_s"]
        else:
            for _ in range(args.steps):
                rate, dt, threads = cpu_port_rate(n_particles, 1, workload="assessor")
                times.append(dt)
            sample = "oracle/port.py PortSwarm per-particle loop (oracle/_ref absent), P=64 x 1 iteration per step"
            kind, label = "port", "restatement"
    ms = 1e3 * float(np.mean(times))
    value = n_particles / (ms / 1e3)
    line = {
        "impl": "reference", "label": label, "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": args.gpus, "steps": args.steps,
        "warmup": args.warmup, "ms_per_step": ms, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
        "dtype": "f32", "data": "synthetic",
        "config": {"workload": WORKLOADS[args.workload].replace("swarm=65,536", f"swarm={args.particles:,}"), "sample": sample,
                   "host": f"{os.cpu_count()} host threads, torch {torch.__version__} CPU"},
        "cpu_baseline": {"value": value, "unit": UNIT, "cores": torch.get_num_threads(), "kind": kind, "label": label, "sample": sample},
        "e2e": {"value": value, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    line.update(extra)
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

def engine_version():
    from psob import _ffi
    return int(_ffi.lib().psob_version())


def make_config3_decoder(seed=3):
    """`VectorQuantizedVAE_GAN.decoder` (src/inverter/utils_vq_vae/util_model.py:283-289): the Generator stack stored under
    `decoder.*`, xavier-uniform conv weights and zero biases (`weights_init`, :39-46)."""
    from src.utils import util_dcgan as M
    torch.manual_seed(seed)
    g = M.Generator(D_LATENT, 1, 64)
    for m in g.modules():
        if isinstance(m, (torch.nn.ConvTranspose2d, torch.nn.Conv2d)):
            torch.nn.init.xavier_uniform_(m.weight.data)
            m.bias.data.fill_(0)
    return {"decoder." + k[len("gen."):]: v for k, v in g.eval().state_dict().items()}


def make_g256(seed=4):
    """EXTENSION generator of BASELINE config 4: the reference's `_block` (ConvT + BN + ReLU) stacked
    4 -> 8 -> ... -> 256; plain torch modules as parameter holders, N(0, 0.02) init like `initialize_weights`."""
    nn = torch.nn
    torch.manual_seed(seed)
    f = 64
    up = lambda ci, co, k, s, p: nn.Sequential(nn.ConvTranspose2d(ci, co, k, s, p), nn.BatchNorm2d(co), nn.ReLU())
    layers, c = [up(D_LATENT, 32 * f, 4, 1, 0)], 32 * f
    for _ in range(5):
        layers.append(up(c, c // 2, 4, 2, 1))
        c //= 2
    layers += [nn.ConvTranspose2d(c, 1, 4, 2, 1), nn.Tanh()]
    gen = nn.Sequential(*layers)
    for m in gen.modules():
        if isinstance(m, (nn.ConvTranspose2d, nn.BatchNorm2d)):
            nn.init.normal_(m.weight.data, 0.0, 0.02)
    return {"gen." + k: v for k, v in gen.eval().state_dict().items()}


def build_swarm(args, P, dev, first, seed, distributed=True, max_iterations=1 << 17):
    """The workload's networks + swarm on `dev` for `P` particles starting at global index `first`."""
    from psob import engine
    extras = {}
    if args.workload in ("dscore", "assessor"):
        gen, disc = make_models()
        gnet = engine.GeneratorNet(gen.state_dict(), P, device=dev)
        if args.workload == "assessor":
            anet = engine.AssessorNet(make_assessor().state_dict(), P, device=dev)
            fitness = engine.ResnetFitness(anet, 0, "optimize_out_training")
            extras["anet"] = anet
        else:
            dnet = engine.DiscriminatorNet(disc.state_dict(), P, device=dev)
            fitness = engine.DScoreFitness(dnet)
            extras["dnet"] = dnet
    else:
        n_pix, tdt = (784, torch.float32) if args.workload == "config3" else (65536, torch.bfloat16)
        sd = make_config3_decoder() if args.workload == "config3" else make_g256()
        gnet = engine.GeneratorNet(sd, P, prefix="decoder" if args.workload == "config3" else "gen", device=dev)
        tg = torch.Generator(device=dev).manual_seed(1000 + first)
        target = (torch.rand(P, n_pix, generator=tg, device=dev) * 2 - 1).to(tdt)
        fitness = engine.ReconFitness(target, w_rec=1.0, eps=0.1)
        extras["target"] = target
    sw = engine.SwarmEngine(P, D_LATENT, generator=gnet, fitness=fitness, w_inertia=0.73, w_cogn=1.496, w_soci=1.496, seed=42,
                            index_offset=first, max_iterations=max_iterations, device=dev, need_image_f32=False,
                            distributed=distributed)
    if seed is not None:
        sw.init_particles(generator=torch.Generator(device=dev).manual_seed(seed))
    return sw, gnet, extras


def shard_parity(args, rank, world, dev):
    """world > 1, before the timed region: 10 iterations of ONE seeded swarm of 8 x 1024 particles sharded over the ranks
    (contiguous blocks, global Philox indices, one record all-gather per iteration) against the same swarm unsharded on rank 0.
    The global-best history, the global-best position and every shard's positions / velocities / personal bests must be
    bit-identical (src/pso/util_pso.py:135-151 applied across ranks)."""
    import torch.distributed as dist
    from psob import dist as pdist
    total, iters = 8 * 1024, 10
    first, count = pdist.shard_range(total, rank, world)
    g = torch.Generator(device=dev).manual_seed(1234)            # the same draws on every rank
    pos = torch.randn(total, D_LATENT, generator=g, device=dev)
    vel = (torch.randn(total, D_LATENT, generator=g, device=dev) - 0.5) / 10
    saved = args.workload
    args.workload = "dscore"
    try:
        sh, _, _ = build_swarm(args, count, dev, first, None, distributed=True, max_iterations=iters)
        sh.inertia_decay = 0.99
        sh.init_particles(pos=pos[first:first + count].contiguous(), vel=vel[first:first + count].contiguous())
        sh.run(iters, use_graph=not args.no_graph)
        torch.cuda.synchronize()
        state = torch.cat([sh.pos, sh.vel, sh.pbest_pos, sh.pbest_val[:, None]], dim=1).contiguous()
        sizes = [pdist.shard_range(total, r, world)[1] for r in range(world)]
        parts = [torch.empty(n, state.shape[1], device=dev) for n in sizes] if rank == 0 else None
        # gather by point-to-point sends (ragged shards)
        if rank == 0:
            parts[0].copy_(state)
            for r in range(1, world):
                dist.recv(parts[r], src=r)
        else:
            dist.send(state, dst=0)
        ok = torch.ones(1, device=dev)
        detail = ""
        if rank == 0:
            whole, _, _ = build_swarm(args, total, dev, 0, None, distributed=False, max_iterations=iters)
            whole.inertia_decay = 0.99
            whole.init_particles(pos=pos, vel=vel)
            whole.run(iters, use_graph=not args.no_graph)
            torch.cuda.synchronize()
            ref = torch.cat([whole.pos, whole.vel, whole.pbest_pos, whole.pbest_val[:, None]], dim=1)
            same_state = torch.equal(torch.cat(parts), ref)
            same_hist = torch.equal(sh.hist, whole.hist) and torch.equal(sh.gbest_pos, whole.gbest_pos) and \
                torch.equal(sh.gbest_val, whole.gbest_val) and sh.iteration_state() == whole.iteration_state()
            improved = int(whole.hist[:iters, 1].sum().item())
            detail = f"state {same_state}, gbest history {same_hist}, {improved} of {iters} iterations improved the global best"
            if not (same_state and same_hist and improved >= 1):
                ok.zero_()
        if rank == 0:
            whole.release_graph()
        dist.broadcast(ok, src=0)
        # every rank must also hold the same global best as rank 0
        gb = torch.cat([sh.gbest_pos, sh.gbest_val]).clone()
        gb0 = gb.clone()
        dist.broadcast(gb0, src=0)
        agree = torch.tensor([1.0 if torch.equal(gb, gb0) else 0.0], device=dev)
        dist.all_reduce(agree, op=dist.ReduceOp.MIN)
        result = bool(ok.item() > 0 and agree.item() > 0)
        sh.release_graph()
        return result, detail
    finally:
        args.workload = saved


def run_psob(args):
    from psob import engine, dist as pdist, _ffi
    import torch.distributed as dist
    engine.require_cuda()
    rank, world, local = pdist.init_from_env("nccl")
    torch.cuda.set_device(local)
    dev = torch.device(f"cuda:{local}")
    strong = args.total_particles > 0
    if strong:
        first, P = pdist.shard_range(args.total_particles, rank, world)
    else:
        P = args.particles
        first = rank * P
    total_particles = args.total_particles if strong else P * world
    assessor = args.workload == "assessor"
    pk = peaks()

    parity, parity_detail = None, ""
    if world > 1:
        parity, parity_detail = shard_parity(args, rank, world, dev)
        torch.cuda.empty_cache()

    sw, gnet, extras = build_swarm(args, P, dev, first, 42 + rank)
    anet, dnet = extras.get("anet"), extras.get("dnet")
    use_graph = not args.no_graph

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def timed_run(n):
        """n whole iterations, state resident in HBM: CUDA events on the launching stream, max over ranks."""
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        barrier()
        e0.record()
        sw.run(n, use_graph=use_graph)
        e1.record()
        barrier()
        ms = torch.tensor([e0.elapsed_time(e1)], device=dev)
        if world > 1:
            dist.all_reduce(ms, op=dist.ReduceOp.MAX)
        return float(ms) / n

    # ---- timed region --------------------------------------------------------------------------------
    sw.run(args.warmup, use_graph=use_graph)
    barrier()
    with ClockSampler(local) as clocks:
        ms_per_step = timed_run(args.steps)
    value = total_particles / (ms_per_step / 1e3)

    # ---- sustained: seconds of back-to-back replays with the clock sampler running ------------------------
    sustained = None
    if args.sustained_s > 0:
        n_sus = max(args.steps, int(args.sustained_s * 1e3 / ms_per_step) + 1)
        with ClockSampler(local) as sclocks:
            sus_ms = timed_run(n_sus)
        sustained = {"seconds": sus_ms * n_sus / 1e3, "steps": n_sus, "ms_per_step": sus_ms, "value": total_particles / (sus_ms / 1e3),
                     "unit": UNIT, "clocks": sclocks.summary()}

    # ---- launches per step: every checked C-ABI call of one eager iteration is one kernel launch ----------
    c0 = _ffi.N_CALLS
    sw.step()
    launches_per_step = _ffi.N_CALLS - c0
    torch.cuda.synchronize()

    # ---- per-kernel breakdown: the engine's own launch list (what step() runs), each entry timed in isolation ----------
    stage_ms = {}
    for name, launch in sw.launch_plan():
        evs = []
        for _ in range(5):
            a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            if name == "gbest":
                barrier()
            a.record()
            launch()
            b.record()
            evs.append((a, b))
        torch.cuda.synchronize()
        stage_ms[name] = stage_ms.get(name, 0.0) + float(np.median([a.elapsed_time(b) for a, b in evs]))
    fit_programs = []
    if assessor:
        fit_programs = [op.prog for op in anet.ops if op.prog is not None]
        stage_ms["assessor_hbm_floor_ms"] = float(sum(p_.meta.get("hbm_bytes", 0) for p_ in fit_programs) * P / (pk["hbm"] * 1e9) * 1e3)
        stage_ms["assessor_tensor_floor_ms"] = float(2.0 * anet.macs_exact * P / (pk["tf_sustained"] * 1e12) * 1e3)
        stage_ms["assessor_total"] = float(sum(v for k, v in stage_ms.items() if k.startswith("a.") or k == "assessor_pack_rescale"))
    elif dnet is not None:
        fit_programs = dnet.programs

    # ---- roofline of the dominant kernel --------------------------------------------------------------------
    progs = {p.name: p for p in list(gnet.programs) + list(fit_programs)}
    top = max((k for k in stage_ms if k in progs), key=lambda k: stage_ms[k])
    flops = 2.0 * progs[top].macs_exact * P
    achieved_tf = flops / (stage_ms[top] * 1e-3) / 1e12
    # DRAM bytes per launch from `ncu --set full` (tools/ncu_summary.py writes profiles/ncu_traffic_latest.json with the engine
    # version and particle count it was captured on); a capture of another engine version / size is refused (null)
    traffic, traffic_note = None, "no ncu capture for this engine version and swarm size"
    tpath = os.path.join(ROOT, "profiles", "ncu_traffic_latest.json")
    if os.path.exists(tpath):
        with open(tpath) as f:
            tj = json.load(f)
        if tj.get("engine_version") == engine_version() and tj.get("particles") == P and top in tj.get("bytes", {}):
            traffic, traffic_note = tj["bytes"][top], f"ncu --set full capture {tj.get('source', '')}"
    roofline = {"kernel": f"pixel_gemm_umma_kernel[{top}]", "bound": "tensor", "achieved": achieved_tf, "peak": pk["tf_burst"],
                "unit": "TFLOP/s", "frac": achieved_tf / pk["tf_burst"], "frac_sustained": achieved_tf / pk["tf_sustained"],
                "peak_kind": "burst (the kernel is timed alone, 5 isolated launches)", "peak_sustained": pk["tf_sustained"],
                "traffic": traffic, "traffic_note": traffic_note, "peak_source": pk["source"],
                "algorithmic_flops_per_launch": flops, "ms_per_launch": stage_ms[top]}
    hbm_bytes = float(progs[top].meta.get("hbm_bytes", 0)) * P
    if hbm_bytes / (pk["hbm"] * 1e9) > flops / (pk["tf_burst"] * 1e12):
        # the dominant layer moves more bytes than it has math for (the assessor's 1x1 convs and shortcut adds):
        # its bound is HBM, algorithmic bytes = every input map once + the output once
        gbs = hbm_bytes / (stage_ms[top] * 1e-3) / 1e9
        roofline = {"kernel": f"pixel_gemm_umma_kernel[{top}]", "bound": "hbm", "achieved": gbs, "peak": pk["hbm"], "unit": "GB/s",
                    "frac": gbs / pk["hbm"], "traffic": traffic, "traffic_note": traffic_note, "peak_source": pk["source"],
                    "algorithmic_bytes_per_launch": hbm_bytes, "algorithmic_flops_per_launch": flops, "ms_per_launch": stage_ms[top]}
    upd_bytes = 5.0 * D_LATENT * 4 * P
    upd_gbs = upd_bytes / (stage_ms["pso_update"] * 1e-3) / 1e9
    roofline_update = {"kernel": "pso_update_kernel", "bound": "hbm", "achieved": upd_gbs, "peak": pk["hbm"], "unit": "GB/s",
                       "frac": upd_gbs / pk["hbm"], "traffic": None, "algorithmic_bytes_per_launch": upd_bytes,
                       "ms_per_launch": stage_ms["pso_update"]}
    net_macs = gnet.macs_exact + (anet.macs_exact if assessor else (dnet.macs_exact if dnet is not None else 0))
    net_flops = 2.0 * net_macs * P
    step_tf = net_flops / (ms_per_step * 1e-3) / 1e12
    gen_ms = sum(stage_ms[p.name] for p in gnet.programs)
    whole_step = {"tflops": step_tf, "frac_burst": step_tf / pk["tf_burst"], "frac_sustained": step_tf / pk["tf_sustained"],
                  "generator_forward_ms": gen_ms, "generator_tflops": 2.0 * gnet.macs_exact * P / (gen_ms * 1e-3) / 1e12}
    if sustained is not None:
        sustained["tflops"] = net_flops / (sustained["ms_per_step"] * 1e-3) / 1e12
        sustained["frac_sustained_peak"] = sustained["tflops"] / pk["tf_sustained"]

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
            if use_graph and sw._graph is not None:
                sw._graph.replay()
            else:
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
    e2e_value = total_particles / (float(e2e_ms) / args.steps / 1e3)

    # ---- the drop-in boundary itself: src.pso.util_pso.Swarm.optimize (N = 1, D-score workload) ----------------
    e2e_dropin = None
    if world == 1 and args.workload == "dscore" and not args.no_dropin:
        e2e_dropin = time_dropin(P, dev)

    # ---- CPU baselines (rank 0, N=1 only), bounded samples ------------------------------------------
    cpu, cpu_same = None, None
    if rank == 0 and world == 1 and not args.no_cpu_baseline:
        torch.set_num_threads(os.cpu_count() or 1)
        try:
            cpu = cpu_reference_config1(n_iterations=10)
        except Exception as e:
            cpu = None
            print(f"cpu_reference_config1 failed: {e!r}", file=sys.stderr)
        if args.workload == "dscore":
            rate, dt, threads = cpu_restatement_dscore(4096, 3)
            cpu_same = {"value": rate, "unit": UNIT, "cores": threads, "kind": "port", "label": "restatement",
                        "sample": f"4096 particles x 3 iterations, batched fp32 torch-CPU restatement of this workload "
                                  f"(oracle/port.py BatchedDScoreSwarm), {dt:.1f} s"}
        elif assessor:
            rate, dt, threads = cpu_port_rate(128, 1, workload="assessor")
            cpu_same = {"value": rate, "unit": UNIT, "cores": threads, "kind": "port", "label": "restatement",
                        "sample": f"128 particles x 1 iteration of the per-particle loop (oracle/port.py PortSwarm), {dt:.1f} s"}
        if cpu is None:
            cpu = cpu_same

    if rank == 0:
        wl = WORKLOADS[args.workload].replace("swarm=65,536", f"swarm={P:,}").replace("swarm=16,384", f"swarm={P:,}").replace("swarm=8,192", f"swarm={P:,}")
        if strong:
            wl += f" -- ONE swarm of {total_particles:,} particles sharded over {world} GPUs"
        ws_gb = P * (2.0 * D_LATENT * 4 * 3 + sum(2.0 * p_.out_pitch for p_ in list(gnet.programs) + list(fit_programs))) / 1e9
        line = {
            "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
            "ms_per_step": ms_per_step, "higher_is_better": True, "scaling": "strong" if strong else "weak", "vs_baseline": None,
            "dtype": "bf16", "data": "synthetic",
            "config": {"workload": wl, "particles_per_gpu": P, "total_particles": total_particles, "latent_dim": D_LATENT,
                       "parallelism": f"dp{world}", "l2_policy": f"inputs_exceed_l2 (per-step working set ~{ws_gb:.1f} GB)",
                       "cuda_graph": bool(use_graph), "state_dtype": "f32", "rng": "philox4x32-10 in-kernel",
                       "engine_version": engine_version()},
            "roofline": roofline, "roofline_update": roofline_update, "whole_step": whole_step,
            "stage_ms": stage_ms, "sustained": sustained,
            "cpu_baseline": cpu, "cpu_baseline_same_workload": cpu_same,
            "e2e": {"value": e2e_value, "unit": UNIT, "h2d_bytes_per_step": 2 * P * D_LATENT * 4,
                    "d2h_bytes_per_step": P * 4 + (D_LATENT + 1) * 4},
            "e2e_dropin": e2e_dropin,
            "gpu_launches": launches_per_step * args.steps, "gpu_launches_per_step": launches_per_step,
            "clocks": clocks.summary(),
        }
        if world > 1:
            line["shard_parity"] = bool(parity)
            line["shard_parity_detail"] = parity_detail
        emit(line)
    if world > 1:
        # Teardown.  The JSON line is out; captured graphs that hold the NCCL all-gather are dropped before the process group
        # (a live one made destroy_process_group wait forever on this stack: torch 2.11 / NCCL 2.28.9), and a watchdog ends the
        # process should the teardown still stall -- a benchmark must never hang on exit.
        code = 0 if parity else 3
        wd = threading.Timer(30.0, lambda: os._exit(code))
        wd.daemon = True
        wd.start()
        sw.release_graph()
        del sw
        import gc
        gc.collect()
        torch.cuda.synchronize()
        dist.barrier()
        dist.destroy_process_group()
        wd.cancel()
        if not parity:
            sys.exit(3)


def time_dropin(P, dev):
    """`src.pso.util_pso.Swarm.optimize` -- the class surface the north_star names -- at the bench's swarm size, once with the
    reference's defaults (early stopping on, diversity metric every iteration: host round trips per iteration) and once as
    a pure device loop (early_stopping=False, no diversity: one CUDA graph replayed n times).  Engine construction (lowering,
    weight upload) is done before the clock starts; each timed call includes init_particles and the final host sync."""
    import contextlib
    import io
    from src.pso import util_pso, util_discovery
    from src.utils import util_dcgan as M
    torch.manual_seed(42)
    gen, dsc = M.Generator(D_LATENT, 1, 64), M.Discriminator(1, 64)
    M.initialize_weights(gen)
    M.initialize_weights(dsc)
    gen, dsc = gen.to(dev).eval(), dsc.to(dev).eval()
    disc = util_discovery.Discovery(iid_class=None, model_gan=gen, model_cnn=dsc, device=str(dev))
    out = {}
    for tag, kw, opt in (("defaults_early_stopping_on", dict(tolerance=0.0), dict(early_stopping=True)),
                         ("device_loop_early_stopping_off", dict(compute_diversity=False, record_history=False), dict(early_stopping=False))):
        n_it = 100
        sw = util_pso.Swarm("/tmp/psob_bench", disc, num_particles=P, n_iterations=n_it, dim_space=D_LATENT, device=str(dev),
                            w_inertia=0.73, w_cogn=1.496, w_soci=1.496, **kw)
        with contextlib.redirect_stdout(io.StringIO()):
            sw.optimize(**opt)                       # builds the engine, captures the graph
            torch.cuda.synchronize()
            t0 = time.perf_counter()
            _, _, _, it = sw.optimize(**opt)
            torch.cuda.synchronize()
            dt = time.perf_counter() - t0
        out[tag] = {"value": P * (it - 1) / dt, "unit": UNIT, "iterations": it - 1, "seconds": dt,
                    "note": "wall clock around Swarm.optimize incl. init_particles; tolerance=0 keeps the early-stopping check "
                            "running without ending the loop" if opt["early_stopping"] else "one CUDA graph replayed per iteration"}
        del sw
    return out


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--impl", default="psob", choices=["psob", "reference"])
    ap.add_argument("--particles", type=int, default=0, help="particles per GPU (weak scaling; default: the workload's BASELINE size)")
    ap.add_argument("--total-particles", type=int, default=0,
                    help="ONE swarm of this many particles sharded over the ranks (strong scaling; BASELINE configs[4]: 1048576)")
    ap.add_argument("--workload", default="dscore", choices=sorted(WORKLOADS),
                    help="dscore: BASELINE configs[1] (headline); assessor: the reference's own ResNet-50 fitness; config3 / config4")
    ap.add_argument("--sustained-s", type=float, default=2.5, help="seconds of back-to-back replays for the `sustained` block (0: off)")
    ap.add_argument("--no-graph", action="store_true")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-dropin", action="store_true")
    args = ap.parse_args()
    if args.particles <= 0:
        args.particles = DEFAULT_PARTICLES[args.workload]
    if args.warmup < 3 and args.impl == "psob":
        args.warmup = 3
    if args.impl == "reference":
        run_reference(args)
    else:
        run_psob(args)


if __name__ == "__main__":
    main()
