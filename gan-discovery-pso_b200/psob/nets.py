"""Weight preparation: reference state_dicts -> BN-folded pixel-GEMM layer programs.

Keys follow the reference modules (src/utils/util_dcgan.py:103-149,
src/inverter/utils_ae/util_inverter.py:164-184,
src/inverter/utils_vq_vae/util_model.py:283-289): `gen.0.0.weight`, `gen.0.1.running_var`,
`disc.2.0.weight`, `enc.3.bias`, `decoder.1.0.weight`, ...
BatchNorm2d in eval() is a per-channel affine of the running statistics, so it folds into
the preceding ConvTranspose2d: w' = w * g/sqrt(var+eps), b' = (b-mean) * g/sqrt(var+eps) + beta.
"""
from __future__ import annotations

from typing import Dict, List, Optional

import os

import numpy as np

from . import lowering as L
from ._knobs import knob

BN_EPS = 1e-5


def _np(v) -> np.ndarray:
    if hasattr(v, "detach"):
        v = v.detach().cpu().numpy()
    return np.asarray(v, dtype=np.float64)


def fold_bn_convt(sd: Dict, conv: str, bn: str):
    """ConvTranspose2d weight [Cin, Cout, k, k] with BN over Cout."""
    w, b = _np(sd[f"{conv}.weight"]), _np(sd[f"{conv}.bias"])
    g, beta = _np(sd[f"{bn}.weight"]), _np(sd[f"{bn}.bias"])
    mean, var = _np(sd[f"{bn}.running_mean"]), _np(sd[f"{bn}.running_var"])
    s = g / np.sqrt(var + BN_EPS)
    return (w * s[None, :, None, None]).astype(np.float32), ((b - mean) * s + beta).astype(np.float32)


def generator_programs(sd: Dict, prefix: str = "gen", m_last: int = 2) -> List[L.LayerProgram]:
    """DCGAN / VQ-VAE-GAN decoder 28x28 generator: z -> 7x7x2f -> 14x14xf -> 28x28x1.  Also the BatchNorm-free
    `VectorQuantizedVAE_MNIST.decoder` (src/inverter/utils_vq_vae/util_model.py:237-244: ConvT / ReLU / ConvT / ReLU /
    ConvT / Tanh under `{prefix}.{0,2,4}`; on a 1x1 input its stride-2 first layer equals the stride-1 form)."""
    if f"{prefix}.0.0.weight" in sd:
        w1, b1 = fold_bn_convt(sd, f"{prefix}.0.0", f"{prefix}.0.1")
        w2, b2 = fold_bn_convt(sd, f"{prefix}.1.0", f"{prefix}.1.1")
        last = f"{prefix}.2"
    else:
        f32 = lambda k: _np(sd[k]).astype(np.float32)
        w1, b1 = f32(f"{prefix}.0.weight"), f32(f"{prefix}.0.bias")
        w2, b2 = f32(f"{prefix}.2.weight"), f32(f"{prefix}.2.bias")
        last = f"{prefix}.4"
    w3 = _np(sd[f"{last}.weight"]).astype(np.float32)
    b3 = _np(sd[f"{last}.bias"]).astype(np.float32)
    assert w1.shape[2] == 7 and w2.shape[2] == 4 and w3.shape[1] == 1, "28x28 single-channel generator expected"
    l1 = L.lower_convt_from_1x1("g.l1", w1, b1, L.ACT_RELU)
    l2 = L.lower_convt_k4s2p1("g.l2", w2, b2, L.ACT_RELU, 7, 7, m=1)
    if L.IMAGE_BAND and w3.shape[0] <= 64:
        l3 = L.lower_convt_image_band("g.l3", w3, b3, L.ACT_TANH)
    else:
        l3 = L.lower_convt_k4s2p1("g.l3", w3, b3, L.ACT_TANH, 14, 14, m=m_last, image_out=True)
    assert l1.out_pitch == l2.in_pitch and l2.out_pitch == l3.in_pitch
    return [l1, l2, l3]


def g256_programs(sd: Dict, prefix: str = "gen") -> List[L.LayerProgram]:
    """EXTENSION generator 1x256x256 (oracle/port.py:G256)."""
    progs = []
    w, b = fold_bn_convt(sd, f"{prefix}.0.0", f"{prefix}.0.1")
    progs.append(L.lower_convt_from_1x1("g256.l0", w, b, L.ACT_RELU))
    h = 4
    i = 1
    while f"{prefix}.{i}.0.weight" in sd:
        w, b = fold_bn_convt(sd, f"{prefix}.{i}.0", f"{prefix}.{i}.1")
        progs.append(L.lower_convt_k4s2p1(f"g256.l{i}", w, b, L.ACT_RELU, h, h, m=1))
        h *= 2
        i += 1
    w = _np(sd[f"{prefix}.{i}.weight"]).astype(np.float32)
    b = _np(sd[f"{prefix}.{i}.bias"]).astype(np.float32)
    if L.IMAGE_BAND and w.shape[0] <= 64:
        # tiled band form: one N = 96 MMA per input pixel scatter-adds its 4x4 patch into an 8-row x 28-column image tile
        progs.append(L.lower_convt_image_band_tiled(f"g256.l{i}", w, b, L.ACT_TANH, h, h))
    else:
        progs.append(L.lower_convt_k4s2p1(f"g256.l{i}", w, b, L.ACT_TANH, h, h, m=2, image_out=True))
    for a, c in zip(progs[:-1], progs[1:]):
        assert a.out_pitch == c.in_pitch, (a.name, a.out_pitch, c.name, c.in_pitch)
    return progs


def _down_programs(sd: Dict, prefix: str, name: str, last_act: int) -> List[L.LayerProgram]:
    f32 = lambda k: _np(sd[k]).astype(np.float32)
    c1 = L.lower_conv_image_k4s2p1(f"{name}.c1", f32(f"{prefix}.0.weight"), f32(f"{prefix}.0.bias"), L.ACT_LEAKY)
    c2 = L.lower_conv_k4s2p1(f"{name}.c2", f32(f"{prefix}.2.0.weight"), f32(f"{prefix}.2.0.bias"), L.ACT_LEAKY, 14, 14)
    c3 = L.lower_conv_full(f"{name}.c3", f32(f"{prefix}.3.weight"), f32(f"{prefix}.3.bias"), last_act)
    assert c1.out_pitch == c2.in_pitch and c2.out_pitch == c3.in_pitch
    return [c1, c2, c3]


def discriminator_programs(sd: Dict, prefix: str = "disc") -> List[L.LayerProgram]:
    """[P,1,28,28] (padded bf16 image) -> sigmoid score fp32 [P,1]."""
    return _down_programs(sd, prefix, "d", L.ACT_SIGMOID)


def head_weights(last: L.LayerProgram) -> np.ndarray:
    """Weights of a `lower_conv_full` layer with one output channel as a vector over its INPUT columns (= the output
    columns of the layer before it): w[pix * cp + ch] = W[0, ch, ky, kx] (src/utils/util_dcgan.py:117), bf16-rounded like
    the layer's own weight pack.  Columns no step reads (channel padding) stay 0."""
    assert last.meta.get("kind") == "conv_full" and last.meta.get("out_c") == 1
    w = np.zeros(last.in_pitch, np.float32)
    for a_col, b_row, _d_col, _flags in np.asarray(last.steps)[:, :4]:
        w[a_col:a_col + L.KCHUNK] = last.wpack[b_row]            # row 0 of the tile = output channel 0
    return w


def encoder_programs(sd: Dict, prefix: str = "enc") -> List[L.LayerProgram]:
    """[P,1,28,28] (padded bf16 image) -> latent fp32 [P,enc_dim]."""
    return _down_programs(sd, prefix, "e", L.ACT_NONE)


def pad_image_bf16(img: np.ndarray) -> np.ndarray:
    """fp32 [P,1,28,28] or [P,784] -> padded layout [P, 31*32] (values bf16-rounded, fp32 storage)."""
    p = img.shape[0]
    out = np.zeros((p, L.IMG_PAD_ROWS, L.IMG_PAD_PITCH), np.float32)
    out[:, 1:29, 1:29] = L.to_bf16_f32(img.reshape(p, 28, 28))
    return out.reshape(p, -1)


def emulate_chain(progs: List[L.LayerProgram], x: np.ndarray, round_out=True):
    """CPU emulation of a whole stack; returns (last bf16 output, last fp32 output)."""
    y, y32 = x, None
    oc = progs[0].meta.get("ones_cols")
    if oc:                                  # the layer's bias rides on two 1.0 columns of the K padding (lower_convt_from_1x1)
        y = np.array(x, dtype=np.float32, copy=True)
        y[:, oc[0]:oc[0] + oc[1]] = 1.0
    for pr in progs:
        y, y32 = L.emulate(pr, y, round_out=round_out)
    return y, y32


# ----------------------------------------------------------------------------------------------
# The reference's assessor: ResNet bottleneck network on the 28x28 image (src/pso/util_cnn.py:81-190)
# ----------------------------------------------------------------------------------------------

def fold_bn_conv(sd: Dict, conv: str, bn: str):
    """Bias-free Conv2d weight [Cout, Cin, k, k] followed by an eval-mode BN over Cout."""
    w = _np(sd[f"{conv}.weight"])
    g, beta = _np(sd[f"{bn}.weight"]), _np(sd[f"{bn}.bias"])
    mean, var = _np(sd[f"{bn}.running_mean"]), _np(sd[f"{bn}.running_var"])
    s = g / np.sqrt(var + BN_EPS)
    return (w * s[:, None, None, None]).astype(np.float32), (beta - mean * s).astype(np.float32)


class AssessorOp:
    """One node of the assessor graph: `kind` in {"layer", "maxpool"}; buffers are named row-major bf16 matrices
    [P, pitch]; a layer reads its whole source row (maps at fixed offsets) and writes `out_elems` elements at
    `dst_off`; `res` names the buffer the epilogue adds before the activation (shortcut mode "epilogue")."""

    def __init__(self, kind, src, dst, prog=None, res=None, pool=None, dst32=None, src_off=0, dst_off=0):
        self.kind, self.src, self.dst, self.prog, self.res, self.pool, self.dst32 = kind, src, dst, prog, res, pool, dst32
        self.src_off, self.dst_off = src_off, dst_off


SHORTCUT_MODE = knob("PSOB_SHORTCUT", "mma")     # "mma": skip connection as MMA steps; "epilogue": residual add in the epilogue


def _assessor_geometry(sd):
    """[(prefix, cin, mid, stride, h_in, h_out, has_projection)] per bottleneck, for a 28x28 input."""
    out, h, c = [], 7, 64
    for li in range(1, 5):
        bi = 0
        while f"layer{li}.{bi}.conv1.weight" in sd:
            pre = f"layer{li}.{bi}"
            mid = int(np.shape(sd[f"{pre}.conv1.weight"])[0])
            stride = 2 if (li > 1 and bi == 0) else 1
            ho = (h + 2 - 3) // stride + 1
            proj = f"{pre}.identity_downsample.0.weight" in sd
            if not proj:
                assert stride == 1 and c == 4 * mid, "an identity shortcut needs matching shapes"
            out.append((pre, c, mid, stride, h, ho, proj))
            h, c = ho, 4 * mid
            bi += 1
    return out, h, c


def assessor_programs(sd: Dict, shortcut: Optional[str] = None):
    """-> (ops, buffers): ops in execution order over named bf16 buffers {name: row pitch in elements}.
    Input buffer "img" = the stem's zero-padded image rows (lowering.pad_image_stem); output "logits" fp32.

    shortcut == "mma" (default): the block input x and the 3x3 conv's output t2 share one row ("cat0"/"cat1":
    [x | t2]), and the third conv accumulates the skip connection as extra MMA steps on the same row -- identity
    blocks for `x += identity`, the projection conv's own weights for `identity_downsample` (src/pso/util_cnn.py:
    182-187) -- so no operand is fetched in the epilogue and the projection needs no layer of its own.
    shortcut == "epilogue": the projection is a separate layer and the third conv adds a residual tensor in its
    epilogue (psob_layer_run_residual)."""
    mode = shortcut or SHORTCUT_MODE
    assert mode in ("mma", "epilogue")
    blocks, h_last, c_last = _assessor_geometry(sd)
    if h_last != 1:
        raise ValueError(f"assessor: final feature map is {h_last}x{h_last}; only inputs that reduce to 1x1 (28x28 images) are "
                         "supported (the global max pool is then the identity)")
    ops, bufs = [], {}
    w, b = fold_bn_conv(sd, "conv1", "bn1")
    stem = L.lower_conv_image_k7s2p3("a.stem", w, b, L.ACT_RELU)
    bufs["img"], bufs["stem"] = stem.in_pitch, stem.out_pitch
    ops.append(AssessorOp("layer", "img", "stem", stem))
    x_max = max([49 * 64] + [ho * ho * 4 * mid for (_, _, mid, _, _, ho, _) in blocks])
    t2_max = max(ho * ho * mid for (_, _, mid, _, _, ho, _) in blocks)
    t2_off = x_max if mode == "mma" else 0
    cat_pitch = x_max + t2_max if mode == "mma" else x_max
    bufs["cat0"] = bufs["cat1"] = cat_pitch
    bufs["t1"] = max(h * h * mid for (_, _, mid, _, h, _, _) in blocks)
    if mode == "epilogue":
        bufs["t2"] = t2_max
        if any(bk[6] for bk in blocks):
            bufs["ds"] = cat_pitch            # a residual operand has the output's row pitch
    ops.append(AssessorOp("maxpool", "stem", "cat0", pool=dict(h=14, w=14, cp=64, k=3, stride=2, pad=1)))
    cur, other = "cat0", "cat1"
    for (pre, c, mid, stride, h, ho, proj) in blocks:
        tag = "a." + pre[5:]
        w1, b1 = fold_bn_conv(sd, f"{pre}.conv1", f"{pre}.bn1")
        w2, b2 = fold_bn_conv(sd, f"{pre}.conv2", f"{pre}.bn2")
        w3, b3 = fold_bn_conv(sd, f"{pre}.conv3", f"{pre}.bn3")
        assert w1.shape[1] == c and w3.shape[0] == 4 * mid
        p1 = L.lower_conv2d(f"{tag}.c1", w1, b1, L.ACT_RELU, h, h, 1, 0, out_pitch=bufs["t1"], in_pitch=cat_pitch)
        ops.append(AssessorOp("layer", cur, "t1", p1))
        if proj:
            ws, bs = fold_bn_conv(sd, f"{pre}.identity_downsample.0", f"{pre}.identity_downsample.1")
        if mode == "mma":
            p2 = L.lower_conv2d(f"{tag}.c2", w2, b2, L.ACT_RELU, h, h, stride, 1, out_base=t2_off, out_pitch=cat_pitch,
                                in_pitch=bufs["t1"])
            ops.append(AssessorOp("layer", "t1", cur, p2))
            sc = ("conv", ws, bs, stride, 0, h, h, L.cpitch(c)) if proj else ("identity", 0)
            p3 = L.lower_conv2d(f"{tag}.c3", w3, b3, L.ACT_RELU, ho, ho, 1, 0, in_base=t2_off, out_pitch=cat_pitch, shortcut=sc,
                                in_pitch=cat_pitch)
            ops.append(AssessorOp("layer", cur, other, p3))
        else:
            p2 = L.lower_conv2d(f"{tag}.c2", w2, b2, L.ACT_RELU, h, h, stride, 1, out_pitch=bufs["t2"], in_pitch=bufs["t1"])
            ops.append(AssessorOp("layer", "t1", "t2", p2))
            res = cur
            if proj:
                pd = L.lower_conv2d(f"{tag}.ds", ws, bs, L.ACT_NONE, h, h, stride, 0, out_pitch=cat_pitch, in_pitch=cat_pitch)
                ops.append(AssessorOp("layer", cur, "ds", pd))
                res = "ds"
            p3 = L.lower_conv2d(f"{tag}.c3", w3, b3, L.ACT_RELU, ho, ho, 1, 0, residual=True, out_pitch=cat_pitch, in_pitch=bufs["t2"])
            ops.append(AssessorOp("layer", "t2", other, p3, res=res))
        cur, other = other, cur
    wf = _np(sd["fc.weight"]).astype(np.float32)
    fc = L.lower_conv_full("a.fc", wf.reshape(wf.shape[0], wf.shape[1], 1, 1), _np(sd["fc.bias"]).astype(np.float32), L.ACT_NONE,
                           in_cp=L.cpitch(c_last), in_pitch=cat_pitch)
    ops.append(AssessorOp("layer", cur, None, fc, dst32="logits"))
    bufs["logits"] = fc.out32_pitch
    return ops, bufs


def emulate_assessor(ops, bufs, img: np.ndarray, round_out=True):
    """img: [P, 28*28] values in the assessor's input range -> logits fp32 [P, n_class] (numpy emulation of the graph)."""
    p = img.shape[0]
    mem = {k: np.zeros((p, n), np.float32) for k, n in bufs.items()}
    mem["img"] = L.pad_image_stem(L.to_bf16_f32(img) if round_out else img.astype(np.float32))
    for op in ops:
        if op.kind == "maxpool":
            q = op.pool
            y = L.maxpool_nhwc(mem[op.src][:, :q["h"] * q["w"] * q["cp"]], q["h"], q["w"], q["cp"], q["k"], q["stride"], q["pad"])
            mem[op.dst][:, :y.shape[1]] = y
            continue
        pr = op.prog
        res = mem[op.res] if op.res else None
        y, y32 = L.emulate(pr, mem[op.src], round_out=round_out, res=res)
        if op.dst:
            b0, n = pr.meta.get("out_base", 0), pr.meta.get("out_elems", pr.out_pitch)
            mem[op.dst][:, b0:b0 + n] = y[:, b0:b0 + n]
        if op.dst32:
            mem[op.dst32] = y32
    return mem["logits"]


# ----------------------------------------------------------------------------------------------
# `VectorQuantizedVAE.decoder` (src/inverter/utils_vq_vae/util_model.py:194-203): two ResBlockBatchNorm(dim) on the 7x7
# latent grid (:151-164), ReLU, ConvT(dim, dim, 4, 2, 1) + BN + ReLU, ConvT(dim, 1, 4, 2, 1) + Tanh.
# ----------------------------------------------------------------------------------------------

def fold_bn_conv_bias(sd: Dict, conv: str, bn: str):
    """Conv2d WITH bias [Cout, Cin, k, k] followed by an eval-mode BN over Cout."""
    w, b = _np(sd[f"{conv}.weight"]), _np(sd[f"{conv}.bias"])
    g, beta = _np(sd[f"{bn}.weight"]), _np(sd[f"{bn}.bias"])
    mean, var = _np(sd[f"{bn}.running_mean"]), _np(sd[f"{bn}.running_var"])
    s = g / np.sqrt(var + BN_EPS)
    return (w * s[:, None, None, None]).astype(np.float32), ((b - mean) * s + beta).astype(np.float32)


def vq_grid_decoder_programs(sd: Dict, prefix: str = "decoder"):
    """-> (ops, buffers, dim): the op graph over named row-major bf16 buffers, like `assessor_programs`.

    The block's first ReLU is in place (`nn.ReLU(True)`), so `x + block(x)` evaluates relu(x) + block(relu(x)); the engine
    therefore keeps every block input ALREADY rectified: the input pack applies the ReLU (psob_pack_grid_nchw), and each
    block's 1x1 conv adds the skip connection as identity ops on the operand ring and ends with ReLU -- which is both the
    next block's in-place ReLU and, after the second block, the decoder's own `nn.ReLU(True)`.
    Buffers: "cat0" / "cat1" = [x | t] (block input and the 3x3 conv's output share a row, as in the assessor),
    "x3" = the rectified output of block 2 in the exact pitch the transposed conv reads, "up" = 14x14 map, image outputs."""
    dim = int(np.shape(sd[f"{prefix}.0.block.1.weight"])[0])
    cp = L.cpitch(dim)
    h = 7
    x_elems = h * h * cp
    cat_pitch = 2 * x_elems
    ops, bufs = [], {"cat0": cat_pitch, "cat1": cat_pitch, "x3": x_elems}
    cur, other = "cat0", "cat1"
    for k in range(2):
        w3, b3 = fold_bn_conv_bias(sd, f"{prefix}.{k}.block.1", f"{prefix}.{k}.block.2")
        w1, b1 = fold_bn_conv_bias(sd, f"{prefix}.{k}.block.4", f"{prefix}.{k}.block.5")
        pa = L.lower_conv2d(f"v.b{k}.c3", w3, b3, L.ACT_RELU, h, h, 1, 1, in_cp=cp, out_base=x_elems, out_pitch=cat_pitch, in_pitch=cat_pitch)
        ops.append(AssessorOp("layer", cur, cur, pa))
        last = k == 1
        pb = L.lower_conv2d(f"v.b{k}.c1", w1, b1, L.ACT_RELU, h, h, 1, 0, in_cp=cp, in_base=x_elems,
                            out_pitch=x_elems if last else cat_pitch, shortcut=("identity", 0), in_pitch=cat_pitch)
        ops.append(AssessorOp("layer", cur, "x3" if last else other, pb))
        cur, other = other, cur
    wu, bu = fold_bn_convt(sd, f"{prefix}.3", f"{prefix}.4")
    up = L.lower_convt_k4s2p1("v.up", wu, bu, L.ACT_RELU, h, h, m=1, in_cp=cp)
    assert up.in_pitch == x_elems
    bufs["up"] = up.out_pitch
    ops.append(AssessorOp("layer", "x3", "up", up))
    wi = _np(sd[f"{prefix}.6.weight"]).astype(np.float32)
    bi = _np(sd[f"{prefix}.6.bias"]).astype(np.float32)
    assert wi.shape[1] == 1, "single-channel image expected"
    img = L.lower_convt_k4s2p1("v.img", wi, bi, L.ACT_TANH, 2 * h, 2 * h, m=2, image_out=True, in_cp=L.cpitch(wu.shape[1]))
    assert img.in_pitch == up.out_pitch
    bufs["img_pad"] = img.out_pitch
    ops.append(AssessorOp("layer", "up", "img_pad", img, dst32="image"))
    bufs["image"] = img.out32_pitch
    return ops, bufs, dim


def emulate_vq_grid_decoder(ops, bufs, dim, z: np.ndarray, round_out=True):
    """z: [P, dim, 7, 7] fp32 -> image fp32 [P, 784] (numpy emulation of the graph; the pack applies ReLU + bf16)."""
    p = z.shape[0]
    cp = L.cpitch(dim)
    mem = {k: np.zeros((p, n), np.float32) for k, n in bufs.items()}
    x = np.zeros((p, 49, cp), np.float32)
    x[:, :, :dim] = np.maximum(z.reshape(p, dim, 49).transpose(0, 2, 1), 0)
    mem["cat0"][:, :49 * cp] = L.to_bf16_f32(x.reshape(p, -1)) if round_out else x.reshape(p, -1)
    for op in ops:
        pr = op.prog
        y, y32 = L.emulate(pr, mem[op.src], round_out=round_out)
        if op.dst and y is not None:
            b0, n = pr.meta.get("out_base", 0), pr.meta.get("out_elems", pr.out_pitch)
            mem[op.dst][:, b0:b0 + n] = y[:, b0:b0 + n]
        if op.dst32:
            mem[op.dst32] = y32
    return mem["image"]
