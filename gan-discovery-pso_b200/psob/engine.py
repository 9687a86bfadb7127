"""Device-side swarm engine: SoA PSO state in HBM, pixel-GEMM networks, one CUDA graph per
iteration.  PyTorch is used for device memory, streams and (multi-GPU) torch.distributed only;
all arithmetic of the hot path runs in libpsob.so kernels through the C ABI (`_ffi`).

Iteration (mirrors Swarm.optimize, src/pso/util_pso.py:167-193, batched over the swarm):

    img          = G(z)                       generator layers (tcgen05 pixel-GEMM)
    fit          = fitness(img)               D-score | reconstruction | assessor callable
    improved,key = pbest/min-loc(fit)
    gbest        = apply(all_gather(candidate record))
    pos,vel,z    = update(...)                Philox r1/r2 in-kernel (or injected)
"""
from __future__ import annotations

import ctypes as C
import math
import os
from typing import Callable, List, Optional

import numpy as np
import torch

from . import _ffi, lowering as L, nets
from ._knobs import knob


def _ptr(t: Optional[torch.Tensor]) -> Optional[int]:
    return None if t is None else t.data_ptr()


def _stream() -> int:
    return torch.cuda.current_stream().cuda_stream


def require_cuda():
    if not torch.cuda.is_available():
        raise _ffi.PsobError("psob needs a CUDA device (B200, sm_100a); there is no CPU fallback")


class Layer:
    """One pixel-GEMM layer program resident on the current device."""

    def __init__(self, prog: L.LayerProgram, in_tiled: bool = False, out_tiled: bool = False, dot_head: bool = False):
        require_cuda()
        self.prog = prog
        self.dot_head = bool(dot_head)
        self.in_tiled, self.out_tiled = bool(in_tiled), bool(out_tiled)
        wbits = np.ascontiguousarray(L.bf16_bits(prog.wpack))
        steps, jobs = L.pack_tables(prog)          # step patterns shared between jobs, relative to a per-job base column
        steps = np.ascontiguousarray(steps, dtype=np.int32)
        jobs = np.ascontiguousarray(jobs, dtype=np.int32)
        chains = np.ascontiguousarray(prog.chains, dtype=np.int32)
        cm = np.ascontiguousarray(prog.col_map, dtype=np.int32)
        cm32 = np.ascontiguousarray(prog.col_map32, dtype=np.int32)
        bias = np.ascontiguousarray(prog.bias, dtype=np.float32)
        d = _ffi.LayerDesc(
            mma_n=prog.mma_n, nt=prog.nt, cc=prog.cc, act=prog.act, in_pitch=prog.in_pitch,
            out_pitch=prog.out_pitch, out32_pitch=prog.out32_pitch, b_resident=int(prog.b_resident),
            vec8=int(prog.vec8), n_steps=int(steps.shape[0]), n_jobs=prog.n_jobs, n_wrows=wbits.shape[0],
            n_bias_rows=bias.shape[0], n_chains=chains.shape[0], chains=chains.ctypes.data, in_tiled=int(in_tiled), out_tiled=int(out_tiled), dot_head=int(dot_head), solo=int(bool(prog.meta.get("residual"))),
            band_pitch=(prog.meta["band"][1] if (prog.meta.get("kind") == "convt_image_band" and knob("PSOB_BAND_EPI", "1") != "0")
                        or prog.meta.get("kind") == "convt_image_band_tiled" else 0),
            band_width=int(prog.meta.get("band_width", 0)), wpack=wbits.ctypes.data, steps=steps.ctypes.data, jobs=jobs.ctypes.data,
            col_map=cm.ctypes.data, col_map32=cm32.ctypes.data, bias=bias.ctypes.data)
        h = C.c_void_p()
        _ffi.check(_ffi.lib().psob_layer_create(C.byref(d), C.byref(h)), f"psob_layer_create({prog.name})")
        self.handle = h

    def run(self, x: torch.Tensor, P: int, y: Optional[torch.Tensor], y32: Optional[torch.Tensor], impl=_ffi.IMPL_TCGEN05):
        assert x.dtype == torch.bfloat16 and x.is_cuda and x.dim() == 2 and x.stride(1) == 1
        _ffi.check(_ffi.lib().psob_layer_run(self.handle, x.data_ptr(), x.stride(0), P, _ptr(y), _ptr(y32), impl, _stream()),
                   f"psob_layer_run({self.prog.name})")

    def __del__(self):
        try:
            if getattr(self, "handle", None):
                _ffi.lib().psob_layer_destroy(self.handle)
                self.handle = None
        except Exception:
            pass


class Network:
    """A stack of layers with its activation workspace for a fixed swarm size."""

    def __init__(self, programs: List[L.LayerProgram], P: int, device=None, impl=_ffi.IMPL_TCGEN05):
        require_cuda()
        self.device = torch.device(device if device is not None else f"cuda:{torch.cuda.current_device()}")
        self.P = P
        self.impl = impl
        self.programs = programs
        self._recon_slots = None
        self.fill_ones = False        # the model nets (generator / discriminator / encoder) write the constant-1 bias columns themselves
        # Internal edges use the tile-major layout [ceil(P/128)][pitch/64][128][64] whenever the producer has the
        # TMA-store epilogue and the consumer reads whole 64-column chunks; the network's own input and its last
        # output stay row-major [P, pitch].  PSOB_TILED=0 keeps everything row-major (A/B measurements).
        tiled_ok = knob("PSOB_TILED", "1") != "0"
        self.edge_tiled = [tiled_ok and L.can_store_tiled(a) and L.can_load_tiled(b) for a, b in zip(programs[:-1], programs[1:])]
        rows = (P + L.TILE_M - 1) // L.TILE_M * L.TILE_M
        with torch.cuda.device(self.device):
            self.layers = [Layer(p, in_tiled=(i > 0 and self.edge_tiled[i - 1]),
                                 out_tiled=(i < len(programs) - 1 and self.edge_tiled[i])) for i, p in enumerate(programs)]
            # zero-initialised so that channel padding and image borders that no job writes stay 0; whole
            # particle tiles are allocated (tile-major buffers are addressed by tile), row-major users see [:P]
            self.acts = [torch.zeros(rows, p.out_pitch, dtype=torch.bfloat16, device=self.device)[:P] if p.out_pitch else None
                         for p in programs]
            self.acts32 = [torch.zeros(P, p.out32_pitch, dtype=torch.float32, device=self.device) if p.out32_pitch else None
                           for p in programs]

    @property
    def in_pitch(self) -> int:
        return self.programs[0].in_pitch

    def launch_plan(self, x: torch.Tensor, write_f32: bool = True, recon=None):
        """The forward pass as a list of (name, launch) pairs: what forward() runs and what bench.py times per stage.

        write_f32=False skips the LAST layer's fp32 store when something else consumes its result (the padded bf16 image
        for the D-score fitness, or `recon`); it is a per-call argument, so a net shared between a swarm and
        `Generator.forward` / `Discovery.particles_to_img_batch` never returns a stale fp32 image.
        recon=(target [P, n_pix] fp32|bf16, part fp32 [recon_slots, P]): the last layer accumulates the per-particle
        sum((target - image)^2) in its epilogue (psob_layer_run_recon, src/pso/util_discovery.py:97)."""
        plan, cur = [], x
        if self.fill_ones and self.programs[0].meta.get("ones_cols"):
            plan.append((self.programs[0].name + ".ones", lambda: _fill_ones(self, x)))
        last = len(self.layers) - 1
        for i, (layer, y, y32) in enumerate(zip(self.layers, self.acts, self.acts32)):
            if i == last and recon is not None:
                target, part = recon
                assert target.is_cuda and target.is_contiguous() and target.dtype in (torch.float32, torch.bfloat16)
                assert part.dtype == torch.float32 and part.shape == (self.recon_slots(), self.P)
                plan.append((layer.prog.name, lambda l=layer, c=cur, y=y, y32=(y32 if write_f32 else None): _ffi.check(
                    _ffi.lib().psob_layer_run_recon(l.handle, c.data_ptr(), c.stride(0), self.P, _ptr(y), _ptr(y32), target.data_ptr(),
                                                    int(target.dtype == torch.bfloat16), target.numel() // self.P, part.data_ptr(),
                                                    _stream()), f"psob_layer_run_recon({l.prog.name})")))
                break
            if i == last and not write_f32 and y is not None:
                y32 = None                      # e.g. D-score fitness only consumes the padded bf16 image
            plan.append((layer.prog.name, lambda l=layer, c=cur, y=y, y32=y32: l.run(c, self.P, y, y32, self.impl)))
            cur = y
        return plan

    def forward(self, x: torch.Tensor, write_f32: bool = True, recon=None):
        """x: bf16 [P, >= in_pitch].  Returns (last bf16 activation or None, last fp32 output or None)."""
        for _, launch in self.launch_plan(x, write_f32=write_f32, recon=recon):
            launch()
        return self.acts[-1], self.acts32[-1]

    def recon_slots(self) -> int:
        """Partial-sum slots per particle of the last layer's reconstruction epilogue (psob_layer_recon_slots)."""
        if self._recon_slots is None:
            n = _ffi.lib().psob_layer_recon_slots(self.layers[-1].handle)
            if n <= 0:
                _ffi.check(n, "psob_layer_recon_slots")
            self._recon_slots = int(n)
        return self._recon_slots

    def activation(self, i: int) -> Optional[torch.Tensor]:
        """Row-major [P, out_pitch] copy of layer i's bf16 output, whatever layout the edge uses (tests, debugging)."""
        y = self.acts[i]
        if y is None or not self.layers[i].out_tiled:
            return y
        return untile(y, self.P)

    @property
    def macs_exact(self) -> int:
        return sum(p.macs_exact for p in self.programs)

    @property
    def macs_issued(self) -> int:
        return sum(p.macs_issued for p in self.programs)


def untile(y: torch.Tensor, P: int) -> torch.Tensor:
    """Tile-major activation buffer (allocated as [rows, pitch], rows a multiple of 128) -> row-major [P, pitch]."""
    pitch = y.shape[1]
    rows = (P + L.TILE_M - 1) // L.TILE_M * L.TILE_M
    flat = torch.as_strided(y, (rows * pitch,), (1,))
    t = flat.view(rows // L.TILE_M, pitch // L.KCHUNK, L.TILE_M, L.KCHUNK).permute(0, 2, 1, 3).reshape(rows, pitch)
    return t[:P].contiguous()


def _sd_numpy(sd):
    return {k: (v.detach().cpu().numpy() if hasattr(v, "detach") else np.asarray(v)) for k, v in sd.items()}


def _fill_ones(net: "Network", x: torch.Tensor):
    """Constant-1 input columns that carry the first layer's bias through the MMA (`ones_cols` of the first program)."""
    oc = net.programs[0].meta.get("ones_cols")
    if oc:
        _ffi.check(_ffi.lib().psob_fill_columns_bf16(x.data_ptr(), net.P, x.stride(0), oc[0], oc[1], 1.0, _stream()),
                   "psob_fill_columns_bf16")


class GeneratorNet(Network):
    """28x28 DCGAN generator / VQ-VAE-GAN decoder (src/utils/util_dcgan.py:128-149)."""

    def __init__(self, state_dict, P, prefix="gen", device=None, impl=_ffi.IMPL_TCGEN05, m_last=None):
        sd = _sd_numpy(state_dict)
        if m_last is None:
            m_last = int(knob("PSOB_M_LAST", "2"))      # image-layer block size (tuning knob)
        if prefix is None:
            prefix = "gen" if any(k.startswith("gen.") for k in sd) else "decoder"
        first = f"{prefix}.0.0.weight" if f"{prefix}.0.0.weight" in sd else f"{prefix}.0.weight"     # BN blocks / BN-free decoder
        self.z_dim = int(sd[first].shape[0])
        if f"{prefix}.3.0.weight" in sd:
            progs = nets.g256_programs(sd, prefix)
        else:
            progs = nets.generator_programs(sd, prefix, m_last=m_last)
        super().__init__(progs, P, device, impl)
        self.fill_ones = True
        self.image_elems = progs[-1].out32_pitch

    def image(self) -> torch.Tensor:
        return self.acts32[-1]

    def image_padded_bf16(self) -> Optional[torch.Tensor]:
        return self.acts[-1]


class DiscriminatorNet(Network):
    """DCGAN discriminator on the padded bf16 image (src/utils/util_dcgan.py:103-125) -> sigmoid score fp32 [P,1]."""

    def __init__(self, state_dict, P, prefix="disc", device=None, impl=_ffi.IMPL_TCGEN05, fused_head=None):
        super().__init__(nets.discriminator_programs(_sd_numpy(state_dict), prefix), P, device, impl)
        self.fill_ones = True
        # Fused head (default on the product path): the hidden conv's epilogue reduces its own output against the
        # final Conv2d(features, 1, k = whole map) -- psob_layer_run_dot + psob_dot_finish -- so the hidden
        # activation is never written and the last layer is not launched.  PSOB_FUSED_HEAD=0 keeps the layer-by-layer
        # path (A/B measurements, per-layer tests).
        if fused_head is None:
            fused_head = knob("PSOB_FUSED_HEAD", "1") != "0"
        self.head = None
        if fused_head and impl == _ffi.IMPL_TCGEN05:
            self.head = self._make_head()

    def _make_head(self):
        hid, last = self.programs[-2], self.programs[-1]
        if last.meta.get("kind") != "conv_full" or last.meta.get("out_c") != 1 or hid.out_pitch != last.in_pitch:
            return None
        if hid.act not in (L.ACT_LEAKY, L.ACT_RELU):
            return None
        lib = _ffi.lib()
        if lib.psob_layer_dot_slots(self.layers[-2].handle) <= 0:
            return None
        # re-create the hidden layer as a dot-only layer (no staging tiles, two epilogue groups); its activation buffer and
        # the score layer's handle are not needed any more
        old = self.layers[-2]
        with torch.cuda.device(self.device):
            self.layers[-2] = Layer(hid, in_tiled=old.in_tiled, out_tiled=False, dot_head=True)
        self.acts[-2] = None
        slots = lib.psob_layer_dot_slots(self.layers[-2].handle)
        assert slots > 0
        w = nets.head_weights(last)                       # fp32 values (bf16-exact) indexed by the hidden layer's output column
        with torch.cuda.device(self.device):
            w_dev = torch.from_numpy(w).to(self.device).to(torch.bfloat16).contiguous()
            part = torch.zeros(slots, self.P, dtype=torch.float32, device=self.device)
        return dict(w=w_dev, part=part, slots=int(slots), bias=float(last.bias[0, 0]), act=int(last.act))

    def launch_plan(self, x: torch.Tensor, finish: bool = True):
        """The forward pass as a list of (name, launch) pairs -- what forward() runs and what bench.py times per stage.
        finish=False leaves the fused head's partial sums for psob_fitness_tail (the swarm's fused iteration)."""
        if not self.head:
            return super().launch_plan(x)
        plan, cur, lib = [], x, _ffi.lib()
        if self.programs[0].meta.get("ones_cols"):
            plan.append((self.programs[0].name + ".ones", lambda: _fill_ones(self, x)))
        n_plain = len(self.layers) - (2 if self.head else 0)
        for layer, y, y32 in list(zip(self.layers, self.acts, self.acts32))[:n_plain]:
            plan.append((layer.prog.name, lambda l=layer, c=cur, y=y, y32=y32: l.run(c, self.P, y, y32, self.impl)))
            cur = y
        if self.head:
            hd, hid = self.head, self.layers[-2]
            plan.append((hid.prog.name, lambda c=cur: _ffi.check(lib.psob_layer_run_dot(
                hid.handle, c.data_ptr(), c.stride(0), self.P, hd["w"].data_ptr(), hd["part"].data_ptr(), _stream()),
                f"psob_layer_run_dot({hid.prog.name})")))
            if finish:
                plan.append((self.programs[-1].name + ".finish", lambda: _ffi.check(lib.psob_dot_finish(
                    hd["part"].data_ptr(), hd["slots"], self.P, hd["bias"], hd["act"], self.acts32[-1].data_ptr(), _stream()),
                    "psob_dot_finish")))
        return plan

    def forward(self, x: torch.Tensor, finish: bool = True):
        for _, launch in self.launch_plan(x, finish=finish):
            launch()
        return (None if self.head else self.acts[-1]), self.acts32[-1]


class EncoderNet(Network):
    """Inverter encoder (src/inverter/utils_ae/util_inverter.py:164-184) -> latent fp32 [P, enc_dim]."""

    def __init__(self, state_dict, P, prefix="enc", device=None, impl=_ffi.IMPL_TCGEN05):
        super().__init__(nets.encoder_programs(_sd_numpy(state_dict), prefix), P, device, impl)
        self.fill_ones = True



class AssessorNet:
    """The reference's ResNet assessor (src/pso/util_cnn.py:81-190, any `layers`) on the pixel-GEMM engine:
    fp32 [P, 784] image (raw generator output, or already in [0, 1]) -> logits fp32 [P, n_class].

    Graph from `nets.assessor_programs`: stem as a banded GEMM on the zero-padded image, 3x3/2 max pool, per
    bottleneck three convs (+ the projection shortcut as its own layer) with the shortcut added in the epilogue of
    the third.  BN folded on the host; activations bf16 NHWC row-major, fp32 accumulation.  The final feature map
    of a 28x28 input is 1x1, so the global max pool is the identity."""

    def __init__(self, state_dict, P, device=None, impl=_ffi.IMPL_TCGEN05, shortcut=None):
        require_cuda()
        self.device = torch.device(device if device is not None else f"cuda:{torch.cuda.current_device()}")
        self.P, self.impl = P, impl
        self.ops, sizes = nets.assessor_programs(_sd_numpy(state_dict), shortcut=shortcut)
        self.n_class = sizes["logits"]
        # Buffers written and read by engine layers only (everything after the max pool) use the tile-major layout
        # [ceil(P/128)][pitch/64][128][64]: every operand tile and epilogue chunk is one contiguous 16 KB block.
        tiled_ok = knob("PSOB_TILED", "1") != "0"
        self.tiled = {k: tiled_ok and k not in ("img", "stem", "logits") for k in sizes}
        rows = (P + L.TILE_M - 1) // L.TILE_M * L.TILE_M
        with torch.cuda.device(self.device):
            self.layers = [Layer(op.prog, in_tiled=self.tiled[op.src], out_tiled=bool(op.dst) and self.tiled[op.dst])
                           if op.kind == "layer" else None for op in self.ops]
            self.buf = {k: torch.zeros(rows, n, dtype=torch.bfloat16, device=self.device)[:P] for k, n in sizes.items() if k != "logits"}
            self.logits = torch.zeros(P, self.n_class, dtype=torch.float32, device=self.device)
            self.rescaled = None

    @property
    def macs_exact(self) -> int:
        return sum(op.prog.macs_exact for op in self.ops if op.prog is not None)

    @property
    def macs_issued(self) -> int:
        return sum(op.prog.macs_issued for op in self.ops if op.prog is not None)

    def pack(self, img: torch.Tensor, rescale: bool, keep_rescaled=False):
        """img fp32 [P, 784] -> the stem's padded bf16 rows; rescale = per-image min-max (src/utils/util_data.py:129-135)."""
        assert img.is_cuda and img.dtype == torch.float32 and img.is_contiguous() and img.numel() == self.P * 784
        if keep_rescaled and self.rescaled is None:
            self.rescaled = torch.empty(self.P, 784, dtype=torch.float32, device=self.device)
        _ffi.check(_ffi.lib().psob_pack_image_stem(img.data_ptr(), self.P, int(bool(rescale)), self.buf["img"].data_ptr(),
                                                   _ptr(self.rescaled) if keep_rescaled else None, _stream()), "psob_pack_image_stem")

    def launch_plan(self, img: Optional[torch.Tensor] = None, rescale=False):
        """(name, launch) pairs of the forward pass: the input pack (when `img` is given) and every node of the graph."""
        lib, plan = _ffi.lib(), []
        if img is not None:
            plan.append(("assessor_pack_rescale", lambda: self.pack(img, rescale)))
        for op, layer in zip(self.ops, self.layers):
            src = self.buf[op.src]
            dst = self.buf[op.dst] if op.dst else None
            if op.kind == "maxpool":
                q = op.pool
                plan.append(("a.maxpool", lambda s_=src, d_=dst, q=q, o=op: _ffi.check(lib.psob_maxpool_nhwc(
                    s_.data_ptr(), s_.stride(0), self.P, q["h"], q["w"], q["cp"], q["k"], q["stride"], q["pad"], d_.data_ptr(),
                    d_.stride(0), int(self.tiled[o.dst]), _stream()), "psob_maxpool_nhwc")))
                continue
            pr = op.prog
            assert dst is None or dst.stride(0) == pr.out_pitch
            res = self.buf[op.res] if op.res else None
            plan.append((pr.name, lambda l=layer, s_=src, d_=dst, o=op, r=res: _ffi.check(lib.psob_layer_run_residual(
                l.handle, s_.data_ptr(), s_.stride(0), self.P, _ptr(d_), _ptr(self.logits if o.dst32 else None), _ptr(r),
                r.stride(0) if r is not None else 0, self.impl, _stream()), f"psob_layer_run({o.prog.name})")))
        return plan

    def run_graph(self):
        for _, launch in self.launch_plan():
            launch()
        return self.logits

    def forward(self, img: torch.Tensor, rescale=False) -> torch.Tensor:
        self.pack(img, rescale)
        return self.run_graph()


class VQGridDecoderNet:
    """`VectorQuantizedVAE.decoder` (src/inverter/utils_vq_vae/util_model.py:194-203) on the pixel-GEMM engine: fp32 latent grid
    [P, dim, 7, 7] -> image fp32 [P, 784] (+ the padded bf16 image).  Graph from `nets.vq_grid_decoder_programs`: per
    ResBlockBatchNorm a 3x3 conv and a 1x1 conv whose accumulator takes the skip connection as identity ops on the operand
    ring (BN folded on the host), then the two transposed convs; all buffers row-major bf16."""

    def __init__(self, state_dict, P, prefix="decoder", device=None, impl=_ffi.IMPL_TCGEN05):
        require_cuda()
        self.device = torch.device(device if device is not None else f"cuda:{torch.cuda.current_device()}")
        self.P, self.impl = P, impl
        self.ops, sizes, self.dim = nets.vq_grid_decoder_programs(_sd_numpy(state_dict), prefix)
        self.cp = L.cpitch(self.dim)
        rows = (P + L.TILE_M - 1) // L.TILE_M * L.TILE_M
        with torch.cuda.device(self.device):
            self.layers = [Layer(op.prog) for op in self.ops]
            self.buf = {k: torch.zeros(rows, n, dtype=torch.bfloat16, device=self.device)[:P] for k, n in sizes.items() if k != "image"}
            self.img32 = torch.zeros(P, sizes["image"], dtype=torch.float32, device=self.device)

    @property
    def macs_exact(self) -> int:
        return sum(op.prog.macs_exact for op in self.ops)

    def launch_plan(self, z: torch.Tensor):
        """z: fp32 [P, dim, 7, 7] contiguous.  (name, launch) pairs: the ReLU + NHWC + bf16 input pack, then the six layers."""
        assert z.is_cuda and z.dtype == torch.float32 and z.is_contiguous() and tuple(z.shape) == (self.P, self.dim, 7, 7)
        lib = _ffi.lib()
        x0 = self.buf["cat0"]
        plan = [("v.pack", lambda: _ffi.check(lib.psob_pack_grid_nchw(z.data_ptr(), self.P, self.dim, 49, self.cp, 1, x0.data_ptr(),
                                                                     x0.stride(0), _stream()), "psob_pack_grid_nchw"))]
        for op, layer in zip(self.ops, self.layers):
            src, dst = self.buf[op.src], self.buf[op.dst]
            plan.append((op.prog.name, lambda l=layer, s_=src, d_=dst, o=op: _ffi.check(lib.psob_layer_run(
                l.handle, s_.data_ptr(), s_.stride(0), self.P, d_.data_ptr(), _ptr(self.img32 if o.dst32 else None), self.impl,
                _stream()), f"psob_layer_run({o.prog.name})")))
        return plan

    def forward(self, z: torch.Tensor) -> torch.Tensor:
        for _, launch in self.launch_plan(z):
            launch()
        return self.img32

    def image(self) -> torch.Tensor:
        return self.img32

    def image_padded_bf16(self) -> torch.Tensor:
        return self.buf["img_pad"]


# ----------------------------------------------------------------------------------------------
# thin functional wrappers over the elementwise entry points (used by the engine and the tests)
# ----------------------------------------------------------------------------------------------

def pack_latent(pos: torch.Tensor, z: torch.Tensor):
    P, D = pos.shape
    _ffi.check(_ffi.lib().psob_pack_latent(pos.data_ptr(), P, D, z.data_ptr(), z.stride(0), _stream()), "psob_pack_latent")


def pack_image(img: torch.Tensor, out: torch.Tensor):
    _ffi.check(_ffi.lib().psob_pack_image(img.data_ptr(), img.shape[0], out.data_ptr(), _stream()), "psob_pack_image")


def rescale_minmax(img: torch.Tensor, out: torch.Tensor):
    P = img.shape[0]
    _ffi.check(_ffi.lib().psob_rescale_minmax(img.data_ptr(), P, img.numel() // P, out.data_ptr(), _stream()), "psob_rescale_minmax")


def mean_pairwise_distance(pos: torch.Tensor) -> float:
    """`Swarm.mse` (src/pso/util_pso.py:76-86) on the device: mean |pos_i - pos_j|_2 over unordered pairs."""
    assert pos.is_cuda and pos.dtype == torch.float32 and pos.dim() == 2 and pos.stride(1) == 1
    P, D = pos.shape
    if P < 2:
        return float("nan")
    total = torch.empty(1, dtype=torch.float64, device=pos.device)
    _ffi.check(_ffi.lib().psob_pairwise_distance_sum(pos.data_ptr(), P, D, pos.stride(0), total.data_ptr(), _stream()),
               "psob_pairwise_distance_sum")
    return float(total.item()) / (P * (P - 1) / 2)


def vq_nearest(x: torch.Tensor, codebook: torch.Tensor, return_distance=False):
    """`vq` (src/inverter/utils_vq_vae/util_function.py:4-23): index of the nearest codebook row for every row
    of x (any leading shape), int64, lowest index on ties."""
    assert x.is_cuda and codebook.is_cuda and x.dtype == torch.float32 and codebook.dtype == torch.float32
    K, D = codebook.shape
    flat = x.reshape(-1, D).contiguous()
    cb = codebook.contiguous()
    idx = torch.empty(flat.shape[0], dtype=torch.int64, device=x.device)
    dist = torch.empty(flat.shape[0], dtype=torch.float32, device=x.device) if return_distance else None
    if flat.shape[0]:
        _ffi.check(_ffi.lib().psob_vq_nearest(flat.data_ptr(), flat.shape[0], D, flat.stride(0), cb.data_ptr(), K, idx.data_ptr(),
                                              _ptr(dist), _stream()), "psob_vq_nearest")
    idx = idx.view(x.shape[:-1])
    return (idx, dist.view(x.shape[:-1])) if return_distance else idx


def score_fitness(score: torch.Tensor, fit: torch.Tensor, mode: str, thr=0.0, eps=0.1):
    if mode not in ("optimize_in_training", "optimize_out_training"):
        raise ValueError(mode)                                  # same error as src/pso/util_discovery.py:80
    m = _ffi.MODE_IN_TRAINING if mode == "optimize_in_training" else _ffi.MODE_OUT_TRAINING
    stride = score.stride(0) if score.dim() > 1 else 1
    _ffi.check(_ffi.lib().psob_score_fitness(score.data_ptr(), stride, score.shape[0], m, thr, eps, fit.data_ptr(), _stream()),
               "psob_score_fitness")


def softmax_select(logits: torch.Tensor, class_idx: int, prob: torch.Tensor):
    P, n = logits.shape
    _ffi.check(_ffi.lib().psob_softmax_select(logits.data_ptr(), n, class_idx, P, prob.data_ptr(), _stream()), "psob_softmax_select")


def recon_fitness(img: torch.Tensor, target: torch.Tensor, fit: torch.Tensor, fit_ass=None, w_ass=1.0, w_rec=1.0, eps=0.1):
    P = img.shape[0]
    n = img.numel() // P
    assert target.numel() == img.numel() and target.dtype in (torch.float32, torch.bfloat16)
    _ffi.check(_ffi.lib().psob_recon_fitness(img.data_ptr(), target.data_ptr(), int(target.dtype == torch.bfloat16), P, n,
                                             _ptr(fit_ass), w_ass, w_rec, eps, fit.data_ptr(), _stream()), "psob_recon_fitness")


def pso_update(pos, vel, pbest_pos, gbest_pos, *, improved=None, r1=None, r2=None, z=None, dyn=None, w=0.5, c_cogn=0.8,
               c_soci=0.9, v_max=math.inf, lo=-math.inf, hi=math.inf, seed=0, index_offset=0, iteration=0):
    P, D = pos.shape
    a = _ffi.UpdateArgs(pos=pos.data_ptr(), vel=vel.data_ptr(), pbest_pos=pbest_pos.data_ptr(), gbest_pos=gbest_pos.data_ptr(),
                        improved=_ptr(improved), r1=_ptr(r1), r2=_ptr(r2), z_bf16=_ptr(z), dyn=_ptr(dyn), P=P, D=D,
                        z_pitch=0 if z is None else z.stride(0), w=w, c_cogn=c_cogn, c_soci=c_soci, v_max=v_max, lo=lo, hi=hi,
                        seed=seed, index_offset=index_offset, iteration=iteration, reserved=0)
    _ffi.check(_ffi.lib().psob_update(C.byref(a), _stream()), "psob_update")


def pbest_minloc(fit, pbest_val, improved, key, index_offset=0):
    _ffi.check(_ffi.lib().psob_pbest_minloc(fit.data_ptr(), pbest_val.data_ptr(), improved.data_ptr(), fit.shape[0],
                                            index_offset, key.data_ptr(), _stream()), "psob_pbest_minloc")


def gbest_candidate(key, pos, pbest_pos, improved, pbest_val, rec, index_offset=0):
    P, D = pos.shape
    _ffi.check(_ffi.lib().psob_gbest_candidate(key.data_ptr(), pos.data_ptr(), pbest_pos.data_ptr(), _ptr(improved),
                                               pbest_val.data_ptr(), P, index_offset, D, rec.data_ptr(), _stream()),
               "psob_gbest_candidate")


def gbest_apply(recs, n_rec, D, gbest_val, gbest_pos, hist=None, dyn=None, inertia_decay=1.0):
    rows = 0 if hist is None else hist.shape[0]
    _ffi.check(_ffi.lib().psob_gbest_apply(recs.data_ptr(), n_rec, D, gbest_val.data_ptr(), gbest_pos.data_ptr(), _ptr(hist),
                                           rows, _ptr(dyn), inertia_decay, _stream()), "psob_gbest_apply")


# ----------------------------------------------------------------------------------------------
# fitness stages
# ----------------------------------------------------------------------------------------------

def _mode_fields(mode: str, thr: float, eps: float) -> dict:
    if mode not in ("optimize_in_training", "optimize_out_training"):
        raise ValueError(mode)
    return dict(mode=_ffi.MODE_IN_TRAINING if mode == "optimize_in_training" else _ffi.MODE_OUT_TRAINING, thr=float(thr), eps=float(eps))


class DScoreFitness:
    """EXTENSION discriminator-score fitness: p = D(G(z)) on the raw image, then the
    clamp/affine of src/pso/util_discovery.py:75-80."""

    def __init__(self, disc: DiscriminatorNet, mode="optimize_out_training", thr=0.0, eps=0.1):
        if mode not in ("optimize_in_training", "optimize_out_training"):
            raise ValueError(mode)
        self.disc, self.mode, self.thr, self.eps = disc, mode, thr, eps
        self.needs_padded_image = True
        self.needs_image_f32 = False

    def __call__(self, gen: GeneratorNet, fit: torch.Tensor):
        _, score = self.disc.forward(gen.image_padded_bf16())
        score_fitness(score, fit, self.mode, self.thr, self.eps)

    def fused_plan(self, gen: GeneratorNet, z: torch.Tensor, write_f32: bool, recon=None):
        """-> ((name, launch) pairs of the networks only, fields telling psob_fitness_tail where the score comes from)."""
        plan = gen.launch_plan(z, write_f32=write_f32, recon=recon) + self.disc.launch_plan(gen.image_padded_bf16(), finish=False)
        d = _mode_fields(self.mode, self.thr, self.eps)
        hd = self.disc.head
        if hd:
            d.update(part=hd["part"].data_ptr(), n_slots=hd["slots"], bias=hd["bias"], act=hd["act"])
        else:
            score = self.disc.acts32[-1]
            d.update(score=score.data_ptr(), score_stride=score.stride(0) if score.dim() > 1 else 1)
        return plan, d


class ReconFitness:
    """Reconstruction / inverter-loss fitness (src/pso/util_discovery.py:84-101):
    w_ass * fitness(pos) + w_rec * mean((target - G(pos))^2) + eps."""

    def __init__(self, target: torch.Tensor, w_rec=1.0, eps=0.1, assessor=None, w_ass=1.0):
        self.target = target.reshape(target.shape[0], -1).contiguous()
        self.w_rec, self.eps, self.assessor, self.w_ass = w_rec, eps, assessor, w_ass
        self.fit_ass = None
        self.part = None
        self._fit_tmp = None
        if assessor is not None:
            self.fit_ass = torch.empty(target.shape[0], dtype=torch.float32, device=target.device)

    @property
    def needs_image_f32(self) -> bool:
        return self.assessor is not None and getattr(self.assessor, "needs_image_f32", True)

    def __call__(self, gen: GeneratorNet, fit: torch.Tensor):
        """Unfused form: reads the materialised fp32 image (tests, A/B measurements, single-particle API)."""
        if self.assessor is not None:
            self.assessor(gen, self.fit_ass)
        recon_fitness(gen.image(), self.target, fit, self.fit_ass, self.w_ass if self.assessor is not None else 0.0,
                      self.w_rec, self.eps)

    def fused_plan(self, gen: GeneratorNet, z: torch.Tensor, write_f32: bool, recon=None):
        """The squared error is accumulated in the generator's last-layer epilogue (never reads the image back);
        psob_fitness_tail adds the slots, divides by the pixel count and forms w_ass * fitness + w_rec * mse + eps.
        Only where the image layer has the band epilogue (one particle row of whole image rows per thread): the 2x2-block
        direct epilogue handles 16 pixels per job and measured SLOWER with the target gather in it (G256's image layer 9.4 ->
        12.2 ms), so such generators keep the materialised image + psob_recon_fitness, whose result enters the tail as a
        ready fitness."""
        assert recon is None
        if not str(gen.programs[-1].meta.get("kind", "")).startswith("convt_image_band"):
            if self.assessor is not None:
                plan = gen.launch_plan(z, write_f32=True) + [("assessor_fitness", lambda: self.assessor(_NoForward(gen), self.fit_ass))]
            else:
                plan = gen.launch_plan(z, write_f32=True)
            if self._fit_tmp is None:
                self._fit_tmp = torch.empty(gen.P, dtype=torch.float32, device=self.target.device)
            plan = plan + [("recon_fitness", lambda: recon_fitness(gen.image(), self.target, self._fit_tmp, self.fit_ass,
                                                                   self.w_ass if self.assessor is not None else 0.0, self.w_rec, self.eps))]
            return plan, dict(fit_in=self._fit_tmp.data_ptr())
        if self.part is None:
            self.part = torch.zeros(gen.recon_slots(), gen.P, dtype=torch.float32, device=self.target.device)
        rc = (self.target, self.part)
        if self.assessor is not None:
            plan, d = self.assessor.fused_plan(gen, z, write_f32 or self.needs_image_f32, recon=rc)
        else:
            plan, d = gen.launch_plan(z, write_f32=write_f32, recon=rc), {}
        d.update(recon_part=self.part.data_ptr(), n_recon_slots=self.part.shape[0], n_pix=self.target.shape[1],
                 w_ass=float(self.w_ass) if self.assessor is not None else 0.0, w_rec=float(self.w_rec), eps_rec=float(self.eps))
        return plan, d


class _NoForward:
    """View of a generator net whose forward has already run (the assessor callables only read its image buffers)."""

    def __init__(self, gen):
        self._gen = gen

    def __getattr__(self, name):
        return getattr(self._gen, name)


class AssessorFitness:
    """The reference's own objective (src/pso/util_discovery.py:52-82) with the assessor CNN as a
    torch callable on the per-image min-max rescaled batch: logits -> softmax pick -> clamp/affine."""

    def __init__(self, model_cnn: Callable, class_index: Optional[int], mode="optimize_out_training", thr=0.0, eps=0.1,
                 image_shape=(1, 28, 28)):
        if mode not in ("optimize_in_training", "optimize_out_training"):
            raise ValueError(mode)
        self.model_cnn, self.class_index, self.mode, self.thr, self.eps = model_cnn, class_index, mode, thr, eps
        self.image_shape = image_shape
        self.rescaled = None
        self.prob = None

    def __call__(self, gen: GeneratorNet, fit: torch.Tensor):
        img = gen.image()
        if self.rescaled is None:
            self.rescaled = torch.empty_like(img)
            self.prob = torch.empty(img.shape[0], dtype=torch.float32, device=img.device)
        rescale_minmax(img, self.rescaled)
        with torch.no_grad():
            logits = self.model_cnn(self.rescaled.view(-1, *self.image_shape)).float().contiguous()
        n_class = logits.shape[1]
        idx = self.class_index if n_class > 2 else 1            # src/pso/util_discovery.py:71-74
        softmax_select(logits, idx, self.prob)
        score_fitness(self.prob, fit, self.mode, self.thr, self.eps)

    needs_image_f32 = True

    def fused_plan(self, gen: GeneratorNet, z: torch.Tensor, write_f32: bool, recon=None):
        img = gen.image()
        if self.rescaled is None:
            self.rescaled = torch.empty_like(img)
            with torch.no_grad():                  # static logits buffer: its address goes into the tail's arguments
                probe = self.model_cnn(self.rescaled[:1].view(-1, *self.image_shape))
            self._logits = torch.zeros(img.shape[0], probe.shape[1], dtype=torch.float32, device=img.device)

        def run_cnn():
            with torch.no_grad():
                self._logits.copy_(self.model_cnn(self.rescaled.view(-1, *self.image_shape)))
        plan = gen.launch_plan(z, write_f32=True, recon=recon) + [("rescale_minmax", lambda: rescale_minmax(img, self.rescaled)),
                                                                  ("model_cnn (torch)", run_cnn)]
        n_class = self._logits.shape[1]
        d = _mode_fields(self.mode, self.thr, self.eps)
        d.update(logits=self._logits.data_ptr(), n_class=n_class, class_idx=self.class_index if n_class > 2 else 1)
        return plan, d


class ResnetFitness:
    """`Discovery.fitness` (src/pso/util_discovery.py:52-82) with the reference's ResNet assessor running on the
    engine (`AssessorNet`): per-image min-max rescale fused into the stem's input pack, logits, softmax pick,
    clamp / affine.  Same arguments as `AssessorFitness`."""

    def __init__(self, net: AssessorNet, class_index: Optional[int], mode="optimize_out_training", thr=0.0, eps=0.1):
        if mode not in ("optimize_in_training", "optimize_out_training"):
            raise ValueError(mode)
        self.net, self.class_index, self.mode, self.thr, self.eps = net, class_index, mode, thr, eps
        self.prob = torch.empty(net.P, dtype=torch.float32, device=net.device)

    def __call__(self, gen: GeneratorNet, fit: torch.Tensor):
        logits = self.net.forward(gen.image(), rescale=True)
        idx = self.class_index if self.net.n_class > 2 else 1      # src/pso/util_discovery.py:71-74
        softmax_select(logits, idx, self.prob)
        score_fitness(self.prob, fit, self.mode, self.thr, self.eps)

    needs_image_f32 = True

    def fused_plan(self, gen: GeneratorNet, z: torch.Tensor, write_f32: bool, recon=None):
        plan = gen.launch_plan(z, write_f32=True, recon=recon) + self.net.launch_plan(gen.image(), rescale=True)
        d = _mode_fields(self.mode, self.thr, self.eps)
        d.update(logits=self.net.logits.data_ptr(), n_class=self.net.n_class, class_idx=self.class_index if self.net.n_class > 2 else 1)
        return plan, d


# ----------------------------------------------------------------------------------------------
# the swarm
# ----------------------------------------------------------------------------------------------

class SwarmEngine:
    """Batched swarm living in HBM.  `fitness(gen, fit_out)` fills fit[P] from the generator's
    outputs; `objective(pos, fit_out)` (generic-objective swarms, src/hands_on/pso.py) may be
    given instead of a generator."""

    def __init__(self, P: int, D: int, *, generator: Optional[GeneratorNet] = None, fitness=None, objective=None,
                 w_inertia=0.5, w_cogn=0.8, w_soci=0.9, v_max=math.inf, lo=-math.inf, hi=math.inf, seed=42,
                 index_offset=0, max_iterations=1024, group=None, device=None, need_image_f32=True, distributed=True,
                 snap_to_codebook: Optional[torch.Tensor] = None):
        require_cuda()
        self.device = torch.device(device if device is not None else f"cuda:{torch.cuda.current_device()}")
        self.P, self.D_logical = P, D
        D = (D + 3) // 4 * 4          # rows are float4-aligned; padding dimensions stay exactly 0 forever
        self.D = D
        self.gen, self.fitness, self.objective = generator, fitness, objective
        # whether the generator's last layer stores its fp32 image: a per-call argument of Network.forward (the net may be
        # shared with Generator.forward / Discovery.*_batch), decided by who consumes the image
        self.write_f32 = bool(need_image_f32) or bool(getattr(fitness, "needs_image_f32", True))
        self.fused_tail = True                   # psob_fitness_tail + psob_gbest_local instead of the 5-launch chain
        self.w0, self.c_cogn, self.c_soci = float(w_inertia), float(w_cogn), float(w_soci)
        self.v_max, self.lo, self.hi = v_max, lo, hi
        self.seed, self.index_offset = int(seed), int(index_offset)
        self.group = group
        self.world = 1
        # distributed=False: a single-rank swarm inside a multi-rank job (e.g. bench.py's unsharded parity run on rank 0)
        if distributed and (group is not None or (torch.distributed.is_available() and torch.distributed.is_initialized())):
            self.world = torch.distributed.get_world_size(group)
        dev = self.device
        f32 = dict(dtype=torch.float32, device=dev)
        self.pos = torch.zeros(P, D, **f32)
        self.vel = torch.zeros(P, D, **f32)
        self.pbest_pos = torch.zeros(P, D, **f32)
        self.pbest_val = torch.full((P,), math.inf, **f32)
        self.fit = torch.zeros(P, **f32)
        self.improved = torch.zeros(P, dtype=torch.uint8, device=dev)
        self.key = torch.full((1,), -1, dtype=torch.int64, device=dev)      # armed (all ones) for psob_fitness_tail
        self.rec = torch.zeros(4 + D, **f32)
        self.recs = torch.zeros(self.world, 4 + D, **f32) if self.world > 1 else self.rec.view(1, -1)
        self.gbest_val = torch.full((1,), math.inf, **f32)
        self.gbest_pos = torch.zeros(D, **f32)
        self.max_iterations = int(max_iterations)
        self.hist = torch.zeros(max_iterations + 1, 2, **f32)
        self._pbest_pending = False               # improved[] not yet folded into pbest_pos by a move
        self._plan = None
        self.dyn = torch.zeros(2, dtype=torch.int32, device=dev)     # psob_iter_state {iteration, w}
        self.z = None
        if generator is not None:
            self.z = torch.zeros(P, generator.in_pitch, dtype=torch.bfloat16, device=dev)
        # VQ-snapped swarm (the quantised-code variant of BASELINE config 3): the particle state stays continuous, but every
        # particle is DECODED at its nearest codebook row -- `vq` (src/inverter/utils_vq_vae/util_function.py:4-23) + embedding
        # lookup between the move and the decoder.  codebook: fp32 [K, D] (e.g. set from PSO positions by `pso_weights`)
        self.codebook = None
        if snap_to_codebook is not None:
            assert generator is not None and snap_to_codebook.shape[1] == self.D_logical
            self.codebook = snap_to_codebook.to(dev, torch.float32).contiguous()
            self.code_idx = torch.zeros(P, dtype=torch.int64, device=dev)
        self.r1 = self.r2 = None
        self.inertia_decay = 1.0
        self._graph = None
        self.reset_iteration()

    # -- state ---------------------------------------------------------------------------------
    def reset_iteration(self):
        st = np.zeros(1, dtype=[("i", np.uint32), ("w", np.float32)])
        st["w"] = self.w0
        self.dyn.copy_(torch.from_numpy(st.view(np.int32).copy()).to(self.device))

    def iteration_state(self):
        raw = self.dyn.cpu().numpy().view(np.uint32)
        return int(raw[0]), float(raw[1:2].view(np.float32)[0])

    def init_particles(self, pos=None, vel=None, generator: Optional[torch.Generator] = None):
        """pos ~ N(0,1), vel = (N(0,1) - 0.5)/10 (src/pso/util_pso.py:25-37) unless given."""
        Dl = self.D_logical
        if pos is None:
            pos = torch.randn(self.P, Dl, generator=generator, device=self.device, dtype=torch.float32)
        if vel is None:
            vel = (torch.randn(self.P, Dl, generator=generator, device=self.device, dtype=torch.float32) - 0.5) / 10
        pos = torch.as_tensor(pos, dtype=torch.float32).to(self.device)
        vel = torch.as_tensor(vel, dtype=torch.float32).to(self.device)
        self.pos.zero_()
        self.vel.zero_()
        self.pos[:, :pos.shape[1]].copy_(pos)
        self.vel[:, :vel.shape[1]].copy_(vel)
        self.pbest_pos.copy_(self.pos)
        self.pbest_val.fill_(math.inf)
        self.gbest_val.fill_(math.inf)
        self.gbest_pos.zero_()
        self.improved.zero_()
        self.key.fill_(-1)
        self._pbest_pending = False
        self.hist.zero_()
        self.reset_iteration()
        if self.z is not None:
            pack_latent(self.pos, self.z)
            self._snap()

    def _snap(self):
        if self.codebook is None:
            return
        lib = _ffi.lib()
        _ffi.check(lib.psob_vq_nearest(self.pos.data_ptr(), self.P, self.D_logical, self.pos.stride(0), self.codebook.data_ptr(),
                                       self.codebook.shape[0], self.code_idx.data_ptr(), None, _stream()), "psob_vq_nearest")
        _ffi.check(lib.psob_gather_pack_latent(self.codebook.data_ptr(), self.code_idx.data_ptr(), self.P, self.D_logical,
                                               self.z.data_ptr(), self.z.stride(0), _stream()), "psob_gather_pack_latent")

    def inject_r(self, r1: Optional[torch.Tensor], r2: Optional[torch.Tensor]):
        """Teacher-forced r1/r2 (parity runs); None restores the in-kernel Philox draws."""
        self.r1 = None if r1 is None else r1.to(self.device, torch.float32).contiguous()
        self.r2 = None if r2 is None else r2.to(self.device, torch.float32).contiguous()

    # -- one iteration -------------------------------------------------------------------------
    def evaluate(self):
        """Networks + fitness into self.fit (the unfused chain: separate finish / score / recon kernels)."""
        if self.gen is not None:
            self.gen.forward(self.z, write_f32=self.write_f32)
            self.fitness(self.gen, self.fit)
        else:
            self.objective(self.pos, self.fit)

    def flush_pbest(self):
        """Fold a pending improved[] into pbest_pos.  The copy normally rides on the next psob_update; this is for callers that
        look at pbest_pos, or evaluate again, between the personal best and the move (never on the graph-captured path)."""
        if self._pbest_pending:
            m = self.improved.bool()
            self.pbest_pos[m] = self.pos[m]
            self.improved.zero_()
            self._pbest_pending = False

    def set_inertia(self, w: float):
        """Overwrite the device-resident inertia weight (Swarm.update_inertia / `swarm.w_inertia = ...`)."""
        it, _ = self.iteration_state()
        st = np.zeros(1, dtype=[("i", np.uint32), ("w", np.float32)])
        st["i"], st["w"] = it, w
        self.dyn.copy_(torch.from_numpy(st.view(np.int32).copy()).to(self.device))

    def _global_best(self):
        lib = _ffi.lib()
        hist_rows = self.hist.shape[0]
        if self.world > 1:
            _ffi.check(lib.psob_gbest_record(self.key.data_ptr(), self.pos.data_ptr(), self.pbest_pos.data_ptr(),
                                             self.improved.data_ptr(), self.pbest_val.data_ptr(), self.P, self.index_offset,
                                             self.D, self.rec.data_ptr(), _stream()), "psob_gbest_record")
            torch.distributed.all_gather_into_tensor(self.recs.view(-1), self.rec, group=self.group)
            gbest_apply(self.recs, self.world, self.D, self.gbest_val, self.gbest_pos, self.hist, self.dyn, self.inertia_decay)
        else:
            _ffi.check(lib.psob_gbest_local(self.key.data_ptr(), self.pos.data_ptr(), self.pbest_pos.data_ptr(),
                                            self.improved.data_ptr(), self.pbest_val.data_ptr(), self.P, self.index_offset, self.D,
                                            self.gbest_val.data_ptr(), self.gbest_pos.data_ptr(), self.hist.data_ptr(), hist_rows,
                                            self.dyn.data_ptr(), self.inertia_decay, _stream()), "psob_gbest_local")

    def personal_and_global_best(self):
        """Personal best + global best from self.fit (unfused: psob_pbest_minloc, then the record / apply kernels)."""
        self.flush_pbest()
        pbest_minloc(self.fit, self.pbest_val, self.improved, self.key, self.index_offset)
        self._pbest_pending = True
        self._global_best()

    def launch_plan(self):
        """The product iteration as (name, launch) pairs, in order: the networks, ONE psob_fitness_tail launch (head finish |
        softmax pick | score -> clamp / affine -> reconstruction term -> personal best -> min-loc key), the global best
        (psob_gbest_local; with ranks: record, all-gather, apply) and the move.  step() runs exactly this list, bench.py times
        its entries one by one.  Built once per engine (all buffers are static)."""
        if self._plan is not None:
            return self._plan
        lib = _ffi.lib()
        if self.gen is not None:
            plan, fields = self.fitness.fused_plan(self.gen, self.z, self.write_f32)
            plan = list(plan)
        else:
            plan, fields = [("objective", lambda: self.objective(self.pos, self.fit))], dict(fit_in=self.fit.data_ptr())
        args = _ffi.FitnessTailArgs(fit=None if fields.get("fit_in") == self.fit.data_ptr() else self.fit.data_ptr(), pbest_val=self.pbest_val.data_ptr(),
                                    improved=self.improved.data_ptr(), P=self.P, index_offset=self.index_offset,
                                    key=self.key.data_ptr(), **fields)

        def tail():
            _ffi.check(lib.psob_fitness_tail(C.byref(args), _stream()), "psob_fitness_tail")
            self._pbest_pending = True
        plan += [("fitness_tail", tail), ("gbest", self._global_best), ("pso_update", self.move)]
        self._plan = plan
        return plan

    def evaluate_and_best(self):
        """The product iteration up to (not including) the move."""
        self.flush_pbest()
        for _, launch in self.launch_plan()[:-1]:
            launch()

    def evaluate_and_personal_best(self):
        """Networks + psob_fitness_tail (`Swarm.update_personal_best`): fitness, personal best and the local min-loc key."""
        self.flush_pbest()
        for _, launch in self.launch_plan()[:-2]:
            launch()

    def global_best(self):
        """`Swarm.update_global_best` after evaluate_and_personal_best()."""
        self._global_best()

    def report(self, first_row: int):
        """(iteration, w, history rows [first_row, iteration) as (values, replaced flags), gbest_pos) in ONE device -> host copy:
        what a host loop with per-iteration decisions (early stopping) needs after every global best."""
        it_guess = min(self.hist.shape[0], first_row + 64)
        packed = torch.cat([self.dyn.view(torch.float32), self.hist[first_row:it_guess].reshape(-1), self.gbest_pos]).cpu().numpy()
        it = int(packed[:1].view(np.uint32)[0])
        w = float(packed[1])
        n = max(0, min(it, it_guess) - first_row)
        rows = packed[2:2 + 2 * (it_guess - first_row)].reshape(-1, 2)[:n]
        if it > it_guess:                      # more than 64 unseen rows (a device loop ran in between): take the slow path once
            vals, rep = self.history()
            return it, w, vals[first_row:], rep[first_row:], packed[-self.D:]
        return it, w, rows[:, 0].copy(), rows[:, 1] > 0.5, packed[-self.D:]

    def move(self):
        pso_update(self.pos, self.vel, self.pbest_pos, self.gbest_pos, improved=self.improved, r1=self.r1, r2=self.r2,
                   z=self.z, dyn=self.dyn, c_cogn=self.c_cogn, c_soci=self.c_soci, v_max=self.v_max, lo=self.lo, hi=self.hi,
                   seed=self.seed, index_offset=self.index_offset)
        self._pbest_pending = False
        self._snap()

    def step(self):
        fused = self.fused_tail and (self.gen is None or hasattr(self.fitness, "fused_plan"))
        if fused:
            self.evaluate_and_best()
        else:
            self.evaluate()
            self.personal_and_global_best()
        self.move()

    # -- graph ---------------------------------------------------------------------------------
    def capture(self):
        """Capture one iteration in a CUDA graph (launch-bound at small P)."""
        it, w = self.iteration_state()
        snap = [t.clone() for t in (self.pos, self.vel, self.pbest_pos, self.pbest_val, self.gbest_val, self.gbest_pos,
                                     self.hist, self.dyn, self.improved, self.key)] + ([self.z.clone()] if self.z is not None else [])
        s = torch.cuda.Stream(device=self.device)
        s.wait_stream(torch.cuda.current_stream())
        with torch.cuda.stream(s):
            self.step()                                    # warm-up: tensor maps, function attributes
        torch.cuda.current_stream().wait_stream(s)
        g = torch.cuda.CUDAGraph()
        with torch.cuda.graph(g):
            self.step()
        self._graph = g
        torch.cuda.synchronize(self.device)
        targets = [self.pos, self.vel, self.pbest_pos, self.pbest_val, self.gbest_val, self.gbest_pos, self.hist, self.dyn,
                   self.improved, self.key] + ([self.z] if self.z is not None else [])
        for t, v in zip(targets, snap):
            t.copy_(v)
        return g

    def release_graph(self):
        """Drop the captured iteration (and the launch list, whose closures keep the engine alive).  Call it before
        torch.distributed.destroy_process_group(): a live CUDA graph that captured the NCCL all-gather keeps the communicator busy
        and the teardown waits on it."""
        self._graph = None
        self._plan = None

    def run(self, n_iterations: int, use_graph=True):
        it, _ = self.iteration_state()
        if it + n_iterations > self.hist.shape[0]:
            raise _ffi.PsobError(f"SwarmEngine.run: {it} + {n_iterations} iterations exceed the history recorder "
                                 f"({self.hist.shape[0]} rows = max_iterations + 1); construct the engine with a larger max_iterations")
        if use_graph and self._graph is None and self.r1 is None:
            self.capture()
        for _ in range(n_iterations):
            if use_graph and self._graph is not None:
                self._graph.replay()
            else:
                self.step()

    def history(self):
        """(g_best_val_dummy, replaced flags) for the iterations run so far."""
        it, _ = self.iteration_state()
        h = self.hist[:min(it, self.hist.shape[0])].cpu().numpy()
        return h[:, 0].copy(), h[:, 1] > 0.5
