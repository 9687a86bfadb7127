"""Host-side lowering of the generator / discriminator / encoder conv stacks to
*pixel-GEMM programs* — the table format executed by the tcgen05 engine in
`csrc/pixel_gemm.cuh` (and, for bring-up, by the SIMT interpreter of the same tables).

Data model (everything the reference computes with `nn.ConvTranspose2d` / `nn.Conv2d`,
src/utils/util_dcgan.py:103-149, becomes chains of 128-particle x 64-channel tiles):

* An activation is a bf16 matrix X[P, pitch]: one row per particle, pixels x channels
  flattened NHWC, channel pitch padded to a multiple of 64 (one 128-byte swizzle row).
* A *step* multiplies one K-chunk: A = X[p0:p0+128, a_col:a_col+64] (TMA tile, M = particles)
  with B = Wpack[b_row:b_row+mma_n, 0:64] (K-major weights) into TMEM columns
  [d_col, d_col+mma_n) of the job's accumulator (overwrite or accumulate).
* A *job* is a list of steps plus an epilogue: column n of the accumulator gets
  bias[n], the layer activation, and is stored at out_base + col_map[n] of the output row,
  if its pixel (n // cc) is enabled in the job's pixel mask.

Geometry (transposed-conv phase structure, borders, paddings) lives entirely in these
tables, which are verified on the CPU against torch's convolutions by `emulate()` before a
GPU ever sees them (tests/test_lowering.py).
"""
from __future__ import annotations

import os
from dataclasses import dataclass, field
from typing import List, Optional, Tuple

import numpy as np

from ._knobs import knob

KCHUNK = 64           # K elements per step: 64 bf16 = one 128 B swizzle row
TILE_M = 128          # particles per tile (UMMA M)
MAX_N = 256           # UMMA N limit / TMEM columns per accumulator buffer

ACT_NONE, ACT_RELU, ACT_LEAKY, ACT_TANH, ACT_SIGMOID = 0, 1, 2, 3, 4
STEP_ACCUM, STEP_B_RESIDENT, STEP_NEXT, STEP_EYE = 1, 2, 4, 8     # NEXT: the op accumulates into the next job of the chain
# op width: an op covers (1 + w) * mma_n accumulator columns and weight rows, w = (flags >> STEP_WIDE_SHIFT) & 3.  Neighbouring output
# pixels of a conv read the same input pixel through neighbouring taps; with the taps' weight tiles stored back to back such ops
# are ONE tcgen05.mma of N = 128 / 192 / 256 instead of 2-4 of N = 64 (whose cost is the ~90-cycle issue floor, not their width).
STEP_WIDE_SHIFT = 8
MERGE_OPS = knob("PSOB_MERGE_OPS", "1") != "0"


def op_width(flags):
    return 1 + ((int(flags) >> STEP_WIDE_SHIFT) & 3)
ROUND_WEIGHTS = True   # tests switch this off to check the geometry exactly in fp32
BIAS_IN_K_IMAGE = knob("PSOB_BIAS_IN_K_IMAGE", "1") != "0"   # ... and of the first discriminator / encoder conv (constant K-chunk)
BIAS_IN_K = knob("PSOB_BIAS_IN_K", "1") != "0"      # first generator layer: bias folded into the K padding of the latent row
# Weight packs up to RESIDENT_SOLO bytes stay in one CTA's shared memory for its lifetime; packs up to
# RESIDENT_PAIR are split over a CTA pair (tcgen05 cta_group::2, csrc/pixel_gemm.cu) when the layer has the
# TMA-store epilogue (64-channel NHWC chunks).  PSOB_PAIR=0 falls back to streaming the weights per step.
RESIDENT_SOLO = 32 * 1024
RESIDENT_PAIR = 256 * 1024 if knob("PSOB_PAIR", "1") != "0" else 0
IMAGE_BAND = knob("PSOB_BAND", "1") != "0"       # band form of the 14x14 -> 28x28 image layer
CHAIN_ROWS = knob("PSOB_CHAIN", "1") != "0"      # share input rows between vertically adjacent jobs (resident layers)
CONV2D_RESIDENT = knob("PSOB_CONV2D_RESIDENT", "1") != "0"   # small general-conv weight packs resident + shared A tiles
CONV2D_RESIDENT_MAX = 80 * 1024
EYE_OPS = knob("PSOB_EYE", "1") != "0"           # identity shortcuts as N = 64 ops on the built-in identity tile
CHAIN_CONVT = knob("PSOB_CHAIN_CONVT", "0") != "0"   # ... also for the transposed convs (measured slower, see below)

# image layouts
IMG_PAD_ROWS, IMG_PAD_PITCH = 31, 32          # bf16 padded 28x28 image: real (y,x) at (y+1,x+1)


def rup(x: int, m: int) -> int:
    return (x + m - 1) // m * m


def cpitch(c: int) -> int:
    """Channel pitch of an NHWC activation with c channels."""
    return rup(c, KCHUNK)


def to_bf16_f32(x: np.ndarray) -> np.ndarray:
    """Round-to-nearest-even fp32 -> bf16, returned as fp32 (numpy has no bf16)."""
    u = np.ascontiguousarray(x, dtype=np.float32).view(np.uint32).astype(np.uint64)
    r = ((u + 0x7FFF + ((u >> 16) & 1)) >> 16) << 16
    return r.astype(np.uint32).view(np.float32).reshape(np.shape(x))


def bf16_bits(x: np.ndarray) -> np.ndarray:
    """fp32 -> bf16 bit pattern (uint16), round-to-nearest-even."""
    u = np.ascontiguousarray(x, dtype=np.float32).view(np.uint32).astype(np.uint64)
    return ((u + 0x7FFF + ((u >> 16) & 1)) >> 16).astype(np.uint16).reshape(np.shape(x))


@dataclass
class LayerProgram:
    name: str
    mma_n: int                    # N of every MMA in this layer (multiple of 16, <= 256)
    nt: int                       # accumulator columns per job (<= 256)
    cc: int                       # accumulator columns per output pixel (mask granularity)
    act: int
    in_pitch: int                 # elements per particle row of the input activation
    out_pitch: int                # bf16 output row pitch (0: no bf16 output)
    out32_pitch: int              # fp32 output row pitch (0: no fp32 output)
    wpack: np.ndarray             # float32 [rows, 64] (already bf16-rounded values)
    b_resident: bool              # whole wpack lives in shared memory for the CTA's lifetime
    steps: np.ndarray             # int32 [n_steps, 4]: a_col, b_row, d_col, flags
    jobs: np.ndarray              # int32 [n_jobs, 8]: step_begin, step_count, out_base, out32_base, mask_lo, mask_hi, 0, 0
    col_map: np.ndarray           # int32 [nt] relative bf16 offsets (-1: not stored)
    col_map32: np.ndarray         # int32 [nt] relative fp32 offsets (-1: not stored)
    bias: np.ndarray              # float32 [n_bias_rows, nt]; jobs[:, 6] selects the row
    vec8: bool                    # col_map is contiguous in aligned groups of 8 (16-byte stores)
    xops: dict = field(default_factory=dict)    # step index -> [(b_row, d_col, flags), ...]: further MMA ops fed by the same A tile
    chains: Optional[np.ndarray] = None          # int32 [n_chains, 2]: first job, job count (default: one job per chain)
    macs_exact: int = 0           # border-aware MACs per particle (reference arithmetic)
    macs_issued: int = 0          # MACs per particle the MMA chain actually executes
    meta: dict = field(default_factory=dict)

    @property
    def n_jobs(self) -> int:
        return int(self.jobs.shape[0])

    @property
    def n_steps(self) -> int:
        return int(self.steps.shape[0])


def can_store_tiled(prog: "LayerProgram") -> bool:
    """The layer's epilogue writes whole, 64-aligned, 64-column bf16 chunks (the TMA-store epilogue of
    csrc/pixel_gemm.cu), so its output can use the tile-major layout."""
    if not (prog.vec8 and prog.nt % KCHUNK == 0 and prog.cc % KCHUNK == 0 and prog.out32_pitch == 0
            and prog.out_pitch % KCHUNK == 0):
        return False
    cm = prog.col_map
    for c in range(0, prog.nt, KCHUNK):
        blk = cm[c:c + KCHUNK]
        if blk[0] < 0:
            if np.any(blk >= 0):
                return False
            continue
        if np.any(np.diff(blk) != 1) or np.any((prog.jobs[:, 2].astype(np.int64) + int(blk[0])) % KCHUNK):
            return False
    return True


def can_load_tiled(prog: "LayerProgram") -> bool:
    """Every step reads a whole 64-column chunk of the input row."""
    return prog.in_pitch % KCHUNK == 0 and bool(np.all(prog.steps[:, 0] % KCHUNK == 0))


def pack_steps(prog: "LayerProgram") -> np.ndarray:
    """Engine step records int32 [n_steps, 8]: a_col, meta, op0..op3, 0, 0 (include/psob.h), one record per step of the program
    (uncompressed; `pack_tables` is what the engine uploads)."""
    out = np.zeros((prog.n_steps, 8), np.int32)

    def enc(b_row, d_col, flags):
        assert b_row % 16 == 0 and b_row // 16 < 65536 and d_col % 2 == 0 and d_col // 2 < 256
        return ((b_row // 16) | ((d_col // 2) << 16) | ((1 << 24) if flags & STEP_NEXT else 0) | ((1 << 25) if flags & STEP_ACCUM else 0) |
                ((1 << 26) if flags & STEP_EYE else 0) | ((op_width(flags) - 1) << 28))       # bit 27: issuer region (engine)

    for i, (a_col, b_row, d_col, flags) in enumerate(prog.steps):
        ops = [(int(b_row), int(d_col), int(flags))] + [tuple(map(int, o)) for o in prog.xops.get(i, [])]
        assert len(ops) <= 4
        has_next = any(f & STEP_NEXT for _, _, f in ops)
        out[i, 0] = a_col
        assert not any(f & STEP_EYE for _, _, f in ops) or len(ops) == 1, "an identity op is the only op of its step"
        out[i, 1] = len(ops) | (8 if flags & STEP_B_RESIDENT else 0) | (16 if has_next else 0) | (32 if flags & STEP_EYE else 0)
        for k, o in enumerate(ops):
            out[i, 2 + k] = enc(*o)
    return out


def pack_jobs(prog: "LayerProgram") -> np.ndarray:
    """Job records with the chain-predecessor flag in bit 0 of word 7 (a_base = 0: uncompressed steps)."""
    jobs = prog.jobs.copy()
    jobs[:, 7] = 0
    for jb, jc in prog.chains:
        jobs[jb + 1:jb + jc, 7] = 1
    return jobs


def pack_tables(prog: "LayerProgram"):
    """(steps int32 [n, 8], jobs int32 [n_jobs, 8]) as uploaded to the engine: step lists are stored RELATIVE to a per-job base
    column and shared between jobs with the same pattern.  A conv layer's jobs differ only by where their window starts (and by a
    handful of border variants), so e.g. G256's 128^2 -> 256^2 image band shrinks from 36 864 step records to a few hundred and fits
    the kernel's shared-memory tables; from global memory every step descriptor costs an L2 round trip (~800 cycles, measured: the
    layer ran at 815 cycles per 100-cycle MMA step).  Job word 7 = chain-predecessor flag | (a_base / 8) << 1."""
    steps = pack_steps(prog)
    jobs = pack_jobs(prog)
    if knob("PSOB_PACK_DEDUP", "1") == "0":          # A/B: one record per step, base column 0
        return steps, jobs
    out_steps, seen = [], {}
    for j in range(jobs.shape[0]):
        sb, sc = int(jobs[j, 0]), int(jobs[j, 1])
        blk = steps[sb:sb + sc].copy()
        base = int(blk[:, 0].min()) // KCHUNK * KCHUNK        # multiple of 64: keeps 16-byte and chunk alignment of every a_col
        blk[:, 0] -= base
        key = blk.tobytes()
        if key not in seen:
            seen[key] = len(out_steps)
            out_steps.extend(blk)
        assert base % 8 == 0 and base // 8 < (1 << 30)
        jobs[j, 0] = seen[key]
        jobs[j, 7] = int(jobs[j, 7]) | ((base // 8) << 1)
    return np.asarray(out_steps, dtype=np.int32).reshape(-1, 8), jobs


def _finish(name, mma_n, nt, cc, act, in_pitch, out_pitch, out32_pitch, tiles, b_resident, steps, jobs,
            col_map, col_map32, bias, macs_exact, meta=None, xops=None, chains=None) -> LayerProgram:
    assert mma_n % 16 == 0 and 16 <= mma_n <= MAX_N and nt <= MAX_N and nt % 16 == 0
    wpack = np.concatenate(tiles, axis=0).astype(np.float32) if tiles else np.zeros((0, KCHUNK), np.float32)
    if ROUND_WEIGHTS:
        wpack = to_bf16_f32(wpack)
    steps = np.asarray(steps, dtype=np.int32).reshape(-1, 4)
    jobs = np.asarray(jobs, dtype=np.int64).reshape(-1, 8)
    jobs = (jobs & 0xFFFFFFFF).astype(np.uint32).view(np.int32)
    col_map = np.asarray(col_map, dtype=np.int32)
    col_map32 = np.asarray(col_map32, dtype=np.int32)
    assert col_map.shape == (nt,) and col_map32.shape == (nt,)
    assert np.all(steps[:, 0] % 8 == 0), "TMA needs 16-byte aligned K-chunk starts"
    vec8 = bool(out_pitch and nt % 8 == 0 and all(
        (col_map[g] >= 0 and col_map[g] % 8 == 0 and np.all(np.diff(col_map[g:g + 8]) == 1)) or
        np.all(col_map[g:g + 8] < 0) for g in range(0, nt, 8)) and cc % 8 == 0)
    bias = np.asarray(bias, dtype=np.float32).reshape(-1, nt)
    assert jobs[:, 6].max(initial=0) < bias.shape[0]
    xops = dict(xops or {})
    if chains is None:
        chains = np.stack([np.arange(jobs.shape[0]), np.ones(jobs.shape[0], np.int64)], axis=1)
    chains = np.asarray(chains, dtype=np.int32).reshape(-1, 2)
    meta = dict(meta or {})
    # compulsory HBM traffic per particle of a layer-by-layer execution: the input row once, the bf16 output once (the fp32
    # image of an image layer is only written on request and is not counted); lowerings that share buffers set it themselves
    meta.setdefault("hbm_bytes", 2 * (int(in_pitch) + int(out_pitch)))
    n_eye = int(np.count_nonzero(steps[:, 3] & STEP_EYE)) if steps.size else 0
    # issued MACs count an op's width (a merged op covers 1-4 x mma_n columns)
    n_ops = sum(op_width(fl) for fl in steps[:, 3] if not fl & STEP_EYE) + sum(op_width(o[2]) for v in xops.values() for o in v) if steps.size else 0
    return LayerProgram(name, mma_n, nt, cc, act, in_pitch, out_pitch, out32_pitch, wpack, b_resident, steps, jobs,
                        col_map, col_map32, bias, vec8, xops, chains, int(macs_exact),
                        n_ops * KCHUNK * mma_n + n_eye * KCHUNK * KCHUNK, meta or {})


def _mask_words(bits: List[bool]) -> Tuple[int, int]:
    v = 0
    for i, b in enumerate(bits):
        if b:
            v |= 1 << i
    assert len(bits) <= 64
    return v & 0xFFFFFFFF, (v >> 32) & 0xFFFFFFFF


# ----------------------------------------------------------------------------------------------
# (a) ConvTranspose2d(z -> C, k, s=1, p=0) on a 1x1 input  == one GEMM  [P,z] x [z, k*k*C]
#     (first generator block, src/utils/util_dcgan.py:135)
# ----------------------------------------------------------------------------------------------

def lower_convt_from_1x1(name, w, b, act, in_pitch=None) -> LayerProgram:
    """w: [Cin, Cout, k, k] fp32 (BN already folded), b: [Cout].  Input X[P, Kp] (z padded to
    a multiple of 64 with zeros); output NHWC [k*k, Cp]."""
    cin, cout, k, _ = w.shape
    kp = rup(cin, KCHUNK)
    in_pitch = in_pitch or kp
    cp = cpitch(cout)
    npix = k * k
    if cp >= MAX_N:
        cc, gpix = MAX_N, 1
    else:
        cc, gpix = cp, MAX_N // cp
    nco = cp // cc
    nt = cc * gpix
    tiles, steps, jobs = [], [], []
    wp = np.zeros((kp, cp, k, k), np.float32)
    wp[:cin, :cout] = w
    bp = np.zeros(cp, np.float32)
    bp[:cout] = b
    meta_extra = {}
    if BIAS_IN_K and kp - cin >= 2:
        # The K padding of the latent row carries the bias: input columns cin, cin + 1 hold 1.0 (written by
        # GeneratorNet.forward / nets.emulate_chain) and their weight rows the bias as a bf16 hi + lo pair (2^-17 relative),
        # so the accumulator already contains it and the epilogue is a bare ReLU + convert -- this layer is bound by its
        # epilogue's instruction chain, not by the two K-steps of its MMA.
        hi = to_bf16_f32(bp)
        wp[cin, :, :, :] = hi[:, None, None]
        wp[cin + 1, :, :, :] = (bp - hi)[:, None, None]
        bp = np.zeros(cp, np.float32)
        meta_extra = dict(ones_cols=(cin, 2))
    for p0 in range(0, npix, gpix):
        for ch in range(nco):
            sb = len(steps)
            for kc in range(kp // KCHUNK):
                t = np.zeros((nt, KCHUNK), np.float32)
                for j in range(gpix):
                    pix = p0 + j
                    if pix < npix:
                        ky, kx = divmod(pix, k)
                        t[j * cc:(j + 1) * cc] = wp[kc * KCHUNK:(kc + 1) * KCHUNK, ch * cc:(ch + 1) * cc, ky, kx].T
                steps.append((kc * KCHUNK, len(tiles) * nt, 0, STEP_ACCUM if kc else 0))
                tiles.append(t)
            lo, hi = _mask_words([p0 + j < npix for j in range(gpix)])
            jobs.append((sb, len(steps) - sb, p0 * cp + ch * cc, 0, lo, hi, ch, 0))
    col_map = [(n // cc) * cp + n % cc for n in range(nt)]
    bias = [[bp[ch * cc + n % cc] for n in range(nt)] for ch in range(nco)]
    return _finish(name, nt, nt, cc, act, in_pitch, npix * cp, 0, tiles, False, steps, jobs, col_map,
                   [-1] * nt, bias, macs_exact=cin * cout * npix,
                   meta=dict(kind="convt1x1", out_hw=(k, k), out_c=cout, out_cp=cp, **meta_extra))


# ----------------------------------------------------------------------------------------------
# (b) ConvTranspose2d(Cin -> Cout, k=4, s=2, p=1): H x W -> 2H x 2W
#     (second / third generator blocks, src/utils/util_dcgan.py:136-137)
#
#     Output rows {2q+1, 2q+2} depend on input rows {q, q+1} only (q = -1 .. H-1), with
#     ky = u%2+2 for the upper input row and ky = u%2 for the lower one.  A block of m x m such
#     pairs (2m x 2m output pixels) reads (m+1) x (m+1) input pixels; every (input pixel,
#     K-chunk) is one step whose B tile holds the taps to all output pixels of the block.
# ----------------------------------------------------------------------------------------------

def _tap(u: int, a: int) -> int:
    """kernel index connecting local input row a to local output row u (or -1)."""
    ky = u + 2 - 2 * a
    return ky if 0 <= ky < 4 else -1


def lower_convt_k4s2p1(name, w, b, act, h, wd, *, m=1, image_out=False, in_cp=None) -> LayerProgram:
    """w: [Cin, Cout, 4, 4], b: [Cout].  Input NHWC [h, wd, Cp_in] -> output NHWC [2h, 2wd, Cp_out],
    or, with image_out (Cout == 1): bf16 padded image (IMG_PAD layout, only for 28x28) and an
    fp32 plain [2h*2wd] image."""
    cin, cout = w.shape[0], w.shape[1]
    cpi = in_cp or cpitch(cin)
    ho, wo = 2 * h, 2 * wd
    npx = (2 * m) ** 2
    if image_out:
        assert cout == 1
        cc, cpo, nco = 1, 1, 1
        nt = rup(npx, 16)
    else:
        cpo = cpitch(cout)
        cc = min(cpo, MAX_N // npx)
        assert cc >= 8 and cpo % cc == 0, "block too large for this channel count"
        nco = cpo // cc
        nt = npx * cc
    wp = np.zeros((rup(cin, KCHUNK), cpo if not image_out else 1, 4, 4), np.float32)
    wp[:cin, :cout] = w
    bp = np.zeros(cpo if not image_out else 1, np.float32)
    bp[:cout] = b
    nkc = rup(cin, KCHUNK) // KCHUNK
    assert nkc * KCHUNK <= cpi

    # position-independent B tiles: index [(a, bb, kc, ch)]
    tiles, tile_of = [], {}
    for a in range(m + 1):
        for bb in range(m + 1):
            for kc in range(nkc):
                for ch in range(nco):
                    t = np.zeros((nt, KCHUNK), np.float32)
                    for u in range(2 * m):
                        ky = _tap(u, a)
                        if ky < 0:
                            continue
                        for v in range(2 * m):
                            kx = _tap(v, bb)
                            if kx < 0:
                                continue
                            n0 = (u * 2 * m + v) * cc
                            t[n0:n0 + cc] = wp[kc * KCHUNK:(kc + 1) * KCHUNK, ch * cc:(ch + 1) * cc, ky, kx].T
                    tile_of[(a, bb, kc, ch)] = len(tiles)
                    tiles.append(t)

    if image_out:
        pad_ok = (ho, wo) == (28, 28)
        row_p, pix_p, base_p = (IMG_PAD_PITCH, 1, IMG_PAD_PITCH + 1) if pad_ok else (0, 0, 0)
        out_pitch = IMG_PAD_ROWS * IMG_PAD_PITCH if pad_ok else 0
        row_f, pix_f = wo, 1
        out32_pitch = ho * wo
    else:
        row_p, pix_p, base_p = wo * cpo, cpo, 0
        out_pitch = ho * wo * cpo
        out32_pitch = 0

    # small weight packs (the Cout = 1 image layer: 9 tiles x 2 KB) stay in shared memory for the CTA's lifetime
    wbytes = len(tiles) * nt * KCHUNK * 2
    resident = wbytes <= RESIDENT_SOLO or (wbytes <= RESIDENT_PAIR and not image_out and cc % 64 == 0 and nt % 64 == 0)
    steps, jobs = [], []
    xops, chains = {}, None
    # measured on B200 (profiles/r01_roles_v4*.log): chaining the 7x7 -> 14x14 layer halves its A loads but costs
    # 25 % more cycles (two N = 256 ops per step alternate between both accumulator buffers and the epilogue loses
    # its overlap), so it stays off by default for transposed convs; PSOB_CHAIN_CONVT=1 switches it on
    chained = resident and CHAIN_ROWS and CHAIN_CONVT and m == 1 and wbytes > RESIDENT_SOLO and h >= 2
    if chained:
        # Vertically adjacent blocks share an input row (block bq reads rows bq and bq+1): the blocks of one block
        # column form a chain; every input row is loaded once per chain and feeds this block with the a = 1 taps and,
        # with STEP_NEXT, the next block with the a = 0 taps (half the A-tile loads; MMA work unchanged).
        chains = []
        for br in range(-1, wd):
            for ch in range(nco):
                chain_begin = len(jobs)
                init = False
                for bq in range(-1, h):
                    sb = len(steps)
                    init_next = False
                    rows = []                       # (input row, a for this job, feeds next job)
                    if bq == h - 1:
                        rows.append((bq, 0, False))                 # the last block re-loads its only row itself
                    if bq + 1 < h:
                        rows.append((bq + 1, 1, bq + 1 < h - 1))
                    for iy, a, dual in rows:
                        for bb in range(2):
                            ix = br + bb
                            if not 0 <= ix < wd:
                                continue
                            for kc in range(nkc):
                                fl = (STEP_ACCUM if init else 0) | STEP_B_RESIDENT
                                init = True
                                if dual:
                                    xops[len(steps)] = [(tile_of[(0, bb, kc, ch)] * nt, 0, STEP_NEXT | (STEP_ACCUM if init_next else 0))]
                                    init_next = True
                                steps.append(((iy * wd + ix) * cpi + kc * KCHUNK, tile_of[(a, bb, kc, ch)] * nt, 0, fl))
                    assert len(steps) > sb and init
                    oy0, ox0 = 2 * bq + 1, 2 * br + 1
                    bits = [0 <= oy0 + u < ho and 0 <= ox0 + v < wo for u in range(2) for v in range(2)]
                    assert any(bits)
                    lo, hi = _mask_words(bits)
                    jobs.append((sb, len(steps) - sb, base_p + oy0 * row_p + ox0 * pix_p + ch * cc, 0, lo, hi, ch, 0))
                    init = init_next
                chains.append((chain_begin, len(jobs) - chain_begin))
    for bq in ([] if chained else range(-1, h, m)):
        for br in range(-1, wd, m):
            for ch in range(nco):
                sb = len(steps)
                for a in range(m + 1):
                    iy = bq + a
                    if not 0 <= iy < h:
                        continue
                    for bb in range(m + 1):
                        ix = br + bb
                        if not 0 <= ix < wd:
                            continue
                        for kc in range(nkc):
                            steps.append(((iy * wd + ix) * cpi + kc * KCHUNK, tile_of[(a, bb, kc, ch)] * nt, 0,
                                          (STEP_ACCUM if len(steps) > sb else 0) | (STEP_B_RESIDENT if resident else 0)))
                oy0, ox0 = 2 * bq + 1, 2 * br + 1
                bits = [0 <= oy0 + u < ho and 0 <= ox0 + v < wo for u in range(2 * m) for v in range(2 * m)]
                if not any(bits):
                    del steps[sb:]
                    continue
                lo, hi = _mask_words(bits)
                base = base_p + oy0 * row_p + ox0 * pix_p + ch * cc
                base32 = oy0 * wo + ox0 if image_out else 0
                jobs.append((sb, len(steps) - sb, base, base32, lo, hi, ch, 0))
    col_map, col_map32 = [], []
    bias = np.zeros((nco, nt), np.float32)
    for n in range(nt):
        pix, c = divmod(n, cc)
        if pix >= npx:
            col_map.append(-1); col_map32.append(-1)
            continue
        u, v = divmod(pix, 2 * m)
        col_map.append(u * row_p + v * pix_p + c if out_pitch else -1)
        col_map32.append(u * row_f + v * pix_f if image_out else -1)
        for ch in range(nco):
            bias[ch, n] = bp[ch * cc + c]
    meta = dict(kind="convt_k4s2p1", out_hw=(ho, wo), out_c=cout, out_cp=cpo, m=m)
    # exact MACs: count connected (input, output) pixel pairs inside the borders
    conn = 0
    for oy in range(ho):
        for ky in range(4):
            num = oy + 1 - ky
            if num % 2 == 0 and 0 <= num // 2 < h:
                conn += 1
    conn_x = 0
    for ox in range(wo):
        for kx in range(4):
            num = ox + 1 - kx
            if num % 2 == 0 and 0 <= num // 2 < wd:
                conn_x += 1
    return _finish(name, nt, nt, cc, act, h * wd * cpi, out_pitch, out32_pitch, tiles, resident, steps, jobs,
                   col_map, col_map32, bias, macs_exact=conn * conn_x * cin * cout, meta=meta, xops=xops, chains=chains)


# ----------------------------------------------------------------------------------------------
# (b') ConvTranspose2d(Cin <= 64 -> 1, k=4, s=2, p=1): 14 x 14 -> the 28 x 28 image, "band" form
#      (last generator block, src/utils/util_dcgan.py:137-138).
#
#      One output channel makes N tiny (16 taps per input pixel) while every A tile still costs 16 KB of
#      shared-memory ingest, so the 2x2-block form (b) is bound by re-reading input pixels (2.25x) and by 8-byte
#      scattered image stores.  Here a job owns a BAND of 8 padded image rows as 8 x 30 accumulator columns
#      (column = band_row * 30 + padded_x): the 4x4 output patch of an input pixel is a strided set of columns,
#      so ONE MMA with a zero-padded, position-independent weight tile (N = 96: rows ky * 30 + kx carry
#      w[:, ky, kx]) scatter-adds the whole patch -- the tensor core does the col2im overlap-add in TMEM.
#      Every input pixel is loaded once per band it touches (238 loads instead of 441) and the band leaves as four
#      128-byte-per-particle TMA stores (two padded image rows each).
# ----------------------------------------------------------------------------------------------

BAND_ROWS, BAND_PITCH, BAND_N = 8, 30, 96


def lower_convt_image_band(name, w, b, act, in_cp=None) -> LayerProgram:
    """w: [Cin, 1, 4, 4], b: [1].  Input NHWC [14, 14, Cp_in] -> bf16 padded image (IMG_PAD layout) and fp32 [784]."""
    cin, cout = w.shape[0], w.shape[1]
    assert cout == 1 and tuple(w.shape[2:]) == (4, 4)
    h = wd = 14
    cpi = in_cp or cpitch(cin)
    nkc = rup(cin, KCHUNK) // KCHUNK
    wp = np.zeros((nkc * KCHUNK, 4, 4), np.float32)
    wp[:cin] = w[:, 0]
    nt = 256
    # weight tiles per K-chunk: full patch, top-cut (ky = 2, 3 land on band rows 0, 1), bottom-cut (ky = 0, 1 land on
    # band rows 6, 7; stored two rows lower so that the 96-column window stays inside the band), and zeros
    tiles, tile_of = [], {}
    for kc in range(nkc):
        wk = wp[kc * KCHUNK:(kc + 1) * KCHUNK]
        full = np.zeros((BAND_N, KCHUNK), np.float32)
        top = np.zeros((BAND_N, KCHUNK), np.float32)
        bot = np.zeros((BAND_N, KCHUNK), np.float32)
        for ky in range(4):
            for kx in range(4):
                full[ky * BAND_PITCH + kx] = wk[:, ky, kx]
                if ky >= 2:
                    top[(ky - 2) * BAND_PITCH + kx] = wk[:, ky, kx]
                else:
                    bot[(ky + 2) * BAND_PITCH + kx] = wk[:, ky, kx]
        for key, t in (("full", full), ("top", top), ("bot", bot)):
            tile_of[(key, kc)] = len(tiles)
            tiles.append(t)
    zero_row = len(tiles) * BAND_N
    tiles.append(np.zeros((BAND_N, KCHUNK), np.float32))
    steps, jobs, xops = [], [], {}
    chain = CHAIN_ROWS
    zero_ops = lambda fl: [(zero_row, 0, fl), (zero_row, 96, fl), (zero_row, 160, fl)]     # three N = 96 clears cover 256 columns
    for band in range(4):
        sb = len(steps)
        # chained (bands 0+1 and 2+3 of a particle tile run back to back on one CTA; chains of all four would leave
        # 512 items for 148 CTAs): input row 4 band + 3 lands on the last two rows of this band AND on the first two rows
        # of the next one, so it is loaded once and feeds both (NEXT op); the next band's accumulator is cleared by NEXT
        # clears on an earlier step of this band.  210 loads instead of 238.
        feeds_next = chain and band in (0, 2)
        fed_by_prev = chain and band in (1, 3)
        rows = range(4 * band if fed_by_prev else 4 * band - 1, 4 * band + 4)
        cleared_next = False
        for iy in rows:
            if not 0 <= iy < h:
                continue
            r0 = 2 * iy - BAND_ROWS * band                  # band row of the patch's first row: -2, 0, 2, 4, 6
            for ix in range(wd):
                for kc in range(nkc):
                    if r0 < 0:
                        key, d_col = "top", 2 * ix
                    elif r0 + 3 < BAND_ROWS:
                        key, d_col = "full", r0 * BAND_PITCH + 2 * ix
                    else:
                        key, d_col = "bot", (r0 - 2) * BAND_PITCH + 2 * ix
                    assert d_col + BAND_N <= nt
                    ops = []
                    if len(steps) == sb and not fed_by_prev:
                        ops += zero_ops(0)          # overlapping windows: no "first touch" per column, clear the band first
                    ops.append((tile_of[(key, kc)] * BAND_N, d_col, STEP_ACCUM))
                    if feeds_next:
                        if key == "bot":            # shared row: top-cut taps into the next band
                            if not cleared_next:
                                # The NEXT clears ride on the LAST step before the shared row: the issuer must own the next
                                # band's accumulator buffer from the first NEXT op on, and that buffer is still being read by
                                # the epilogue of the job before this one -- acquiring it on the first step of the band would
                                # serialise that epilogue (the longest of all layers: tanh + masks) with this band's MMAs.
                                prev = len(steps) - 1
                                assert prev >= sb and prev not in xops
                                xops[prev] = zero_ops(STEP_NEXT)
                                cleared_next = True
                            ops.append((tile_of[("top", kc)] * BAND_N, 2 * ix, STEP_ACCUM | STEP_NEXT))
                    assert len(ops) <= 4
                    b_row, dc, fl = ops[0]
                    if len(ops) > 1:
                        xops[len(steps)] = ops[1:]
                    steps.append(((iy * wd + ix) * cpi + kc * KCHUNK, b_row, dc, fl | STEP_B_RESIDENT))
        assert cleared_next == feeds_next
        py0 = BAND_ROWS * band
        lo, hi = _mask_words([1 <= py0 + r <= 28 for r in range(BAND_ROWS)])
        jobs.append((sb, len(steps) - sb, py0 * IMG_PAD_PITCH, (py0 - 1) * 28, lo, hi, 0, 0))
    chains = [(0, 2), (2, 2)] if chain else None
    col_map, col_map32 = [], []
    for n in range(nt):
        r, px = divmod(n, BAND_PITCH)
        ok = r < BAND_ROWS and 1 <= px <= 28
        col_map.append(r * IMG_PAD_PITCH + px if ok else -1)
        col_map32.append(r * 28 + px - 1 if ok else -1)
    bias = np.full(nt, float(np.asarray(b).reshape(-1)[0]), np.float32)
    conn = sum(1 for o in range(28) for k in range(4) if (o + 1 - k) % 2 == 0 and 0 <= (o + 1 - k) // 2 < 14)
    return _finish(name, BAND_N, nt, BAND_PITCH, act, h * wd * cpi, IMG_PAD_ROWS * IMG_PAD_PITCH, 28 * 28, tiles, True, steps, jobs,
                   col_map, col_map32, bias, macs_exact=conn * conn * cin, xops=xops, chains=chains,
                   meta=dict(kind="convt_image_band", out_hw=(28, 28), out_c=1, out_cp=1, band=(BAND_ROWS, BAND_PITCH)))


# ----------------------------------------------------------------------------------------------
# (b'') The band form for WIDE images (G256's 128 x 128 -> 256 x 256 image layer): a job owns a TILE of 8 padded image rows x
#       28 image columns (+ one halo column on each side = the same 8 x 30 accumulator as (b')), so that ONE N = 96 MMA per
#       input pixel still scatter-adds its whole 4 x 4 patch.  Input pixels left / right of the tile contribute only their two
#       nearest taps (cut weight tiles "L" / "R", like "top" / "bot" for the rows); halo columns collect partial sums that
#       the epilogue discards.  There is no bf16 padded image: the epilogue (csrc/pixel_gemm.cu, band == 2) stores fp32 pixels
#       and / or accumulates the reconstruction loss, per particle row, straight from the accumulator.
#       Job mask: bits 0-7 valid padded rows of the band, bits 8-15 valid image columns of the tile (28, or less at the
#       right border).
# ----------------------------------------------------------------------------------------------

BAND_TILE_W = 28


def lower_convt_image_band_tiled(name, w, b, act, h, wd, in_cp=None) -> LayerProgram:
    """w: [Cin, 1, 4, 4], b: [1].  Input NHWC [h, wd, Cp_in] -> fp32 image [2h * 2wd] (no bf16 output)."""
    cin, cout = w.shape[0], w.shape[1]
    assert cout == 1 and tuple(w.shape[2:]) == (4, 4)
    cpi = in_cp or cpitch(cin)
    nkc = rup(cin, KCHUNK) // KCHUNK
    ho, wo = 2 * h, 2 * wd
    assert wo % 4 == 0
    wp = np.zeros((nkc * KCHUNK, 4, 4), np.float32)
    wp[:cin] = w[:, 0]
    nt = 256
    rowpos = {"full": lambda ky: ky, "top": lambda ky: ky - 2 if ky >= 2 else None, "bot": lambda ky: ky + 2 if ky < 2 else None}
    colpos = {"mid": lambda kx: kx, "L": lambda kx: kx - 2 if kx >= 2 else None, "R": lambda kx: kx if kx < 2 else None}
    tiles, tile_of = [], {}
    for kc in range(nkc):
        wk = wp[kc * KCHUNK:(kc + 1) * KCHUNK]
        for rk, rf in rowpos.items():
            for ck, cf in colpos.items():
                t = np.zeros((BAND_N, KCHUNK), np.float32)
                for ky in range(4):
                    for kx in range(4):
                        r, c = rf(ky), cf(kx)
                        if r is not None and c is not None:
                            t[r * BAND_PITCH + c] = wk[:, ky, kx]
                tile_of[(rk, ck, kc)] = len(tiles)
                tiles.append(t)
    zero_row = len(tiles) * BAND_N
    tiles.append(np.zeros((BAND_N, KCHUNK), np.float32))
    n_bands = (ho + 2 + BAND_ROWS - 1) // BAND_ROWS
    n_ct = (wo + BAND_TILE_W - 1) // BAND_TILE_W
    steps, jobs, xops = [], [], {}
    for band in range(n_bands):
        for ct in range(n_ct):
            sb = len(steps)
            for iy in range(4 * band - 1, 4 * band + 4):
                if not 0 <= iy < h:
                    continue
                r0 = 2 * iy - BAND_ROWS * band                  # band row of the patch's first row: -2, 0, 2, 4, 6
                rk, rbase = ("top", 0) if r0 < 0 else (("full", r0) if r0 + 3 < BAND_ROWS else ("bot", r0 - 2))
                for ix in range(14 * ct - 1, 14 * ct + 15):
                    if not 0 <= ix < wd:
                        continue
                    c0 = 2 * (ix - 14 * ct)                     # padded tile column of the patch's first column: -2 .. 28
                    ck, cbase = ("L", 0) if c0 < 0 else (("mid", c0) if c0 + 3 < BAND_PITCH else ("R", c0))
                    for kc in range(nkc):
                        d_col = rbase * BAND_PITCH + cbase
                        assert d_col % 2 == 0 and d_col + BAND_N <= nt
                        ops = []
                        if len(steps) == sb:
                            ops += [(zero_row, 0, 0), (zero_row, 96, 0), (zero_row, 160, 0)]      # three N = 96 clears cover 256 columns
                        ops.append((tile_of[(rk, ck, kc)] * BAND_N, d_col, STEP_ACCUM))
                        b_row, dc, fl = ops[0]
                        if len(ops) > 1:
                            xops[len(steps)] = ops[1:]
                        steps.append(((iy * wd + ix) * cpi + kc * KCHUNK, b_row, dc, fl | STEP_B_RESIDENT))
            assert len(steps) > sb
            py0 = BAND_ROWS * band
            n_cols = min(BAND_TILE_W, wo - BAND_TILE_W * ct)
            bits = [1 <= py0 + r <= ho for r in range(BAND_ROWS)] + [bool((n_cols >> i) & 1) for i in range(8)]
            lo, hi = _mask_words(bits)
            jobs.append((sb, len(steps) - sb, 0, (py0 - 1) * wo + BAND_TILE_W * ct, lo, hi, 0, 0))
    col_map32 = []
    for n in range(nt):
        r, px = divmod(n, BAND_PITCH)
        col_map32.append(r * wo + px - 1 if (r < BAND_ROWS and 1 <= px <= BAND_TILE_W) else -1)
    bias = np.full(nt, float(np.asarray(b).reshape(-1)[0]), np.float32)
    conn_y = sum(1 for o in range(ho) for k in range(4) if (o + 1 - k) % 2 == 0 and 0 <= (o + 1 - k) // 2 < h)
    conn_x = sum(1 for o in range(wo) for k in range(4) if (o + 1 - k) % 2 == 0 and 0 <= (o + 1 - k) // 2 < wd)
    return _finish(name, BAND_N, nt, BAND_PITCH, act, h * wd * cpi, 0, ho * wo, tiles, True, steps, jobs,
                   [-1] * nt, col_map32, bias, macs_exact=conn_y * conn_x * cin, xops=xops,
                   meta=dict(kind="convt_image_band_tiled", out_hw=(ho, wo), out_c=1, out_cp=1, band=(BAND_ROWS, BAND_PITCH),
                             band_width=wo))


# ----------------------------------------------------------------------------------------------
# (c) Conv2d(Cin -> Cout, k=4, s=2, p=1): H x W -> H/2 x W/2  (second discriminator / encoder
#     conv, src/utils/util_dcgan.py:112, 116-121).  One accumulator of mma_n = Cp_out columns
#     per output pixel, g pixels per job; taps whose input pixel is outside are dropped.
# ----------------------------------------------------------------------------------------------

def lower_conv_k4s2p1(name, w, b, act, h, wd, in_cp=None) -> LayerProgram:
    cout, cin = w.shape[0], w.shape[1]
    cpi = in_cp or cpitch(cin)
    cpo = cpitch(cout)
    assert cpo <= MAX_N
    ho, wo = h // 2, wd // 2
    g = MAX_N // cpo
    nt = g * cpo
    nkc = rup(cin, KCHUNK) // KCHUNK
    wp = np.zeros((cpo, rup(cin, KCHUNK), 4, 4), np.float32)
    wp[:cout, :cin] = w
    bp = np.zeros(cpo, np.float32)
    bp[:cout] = b
    tiles, tile_of = [], {}
    for ky in range(4):
        for kx in range(4):
            for kc in range(nkc):
                tile_of[(ky, kx, kc)] = len(tiles)
                tiles.append(wp[:, kc * KCHUNK:(kc + 1) * KCHUNK, ky, kx].copy())
    wbytes = len(tiles) * cpo * KCHUNK * 2
    resident = RESIDENT_SOLO < wbytes <= RESIDENT_PAIR and cpo % 64 == 0       # CTA-pair mode only (small packs stream fine)
    if resident and CHAIN_ROWS:
        return _lower_conv_k4s2p1_chained(name, wp, bp, act, h, wd, cin, cout, cpi, cpo, g, nt, nkc, tiles, tile_of)
    steps, jobs = [], []
    macs = 0
    for oy in range(ho):
        for ox0 in range(0, wo, g):
            sb = len(steps)
            for j in range(g):
                ox = ox0 + j
                if ox >= wo:
                    continue
                first = True
                for ky in range(4):
                    iy = 2 * oy - 1 + ky
                    if not 0 <= iy < h:
                        continue
                    for kx in range(4):
                        ix = 2 * ox - 1 + kx
                        if not 0 <= ix < wd:
                            continue
                        macs += cin * cout
                        for kc in range(nkc):
                            steps.append(((iy * wd + ix) * cpi + kc * KCHUNK, tile_of[(ky, kx, kc)] * cpo, j * cpo,
                                          (0 if first else STEP_ACCUM) | (STEP_B_RESIDENT if resident else 0)))
                            first = False
            lo, hi = _mask_words([ox0 + j < wo for j in range(g)])
            jobs.append((sb, len(steps) - sb, (oy * wo + ox0) * cpo, 0, lo, hi, 0, 0))
    col_map = list(range(nt))
    bias = [bp[n % cpo] for n in range(nt)]
    return _finish(name, cpo, nt, cpo, act, h * wd * cpi, ho * wo * cpo, 0, tiles, resident, steps, jobs, col_map,
                   [-1] * nt, bias, macs_exact=macs, meta=dict(kind="conv_k4s2p1", out_hw=(ho, wo), out_c=cout, out_cp=cpo))


def _lower_conv_k4s2p1_chained(name, wp, bp, act, h, wd, cin, cout, cpi, cpo, g, nt, nkc, tiles, tile_of) -> LayerProgram:
    """Resident-weight form of (c): the jobs of one output-column group (g pixels wide) form a CHAIN over the
    output rows.  Output rows oy and oy+1 share input rows 2oy+1 and 2oy+2, and the g pixels of a job share input
    columns, so every A tile (one input pixel, one K-chunk) is loaded ONCE per chain and feeds up to 2g MMA ops:
    the taps of this job's pixels and, with STEP_NEXT, the taps (ky - 2) of the next job's pixels.  A tile loads per
    output pixel drop from ~14 to ~6 for the 14x14 -> 7x7 layer; the MMA work is unchanged."""
    ho, wo = h // 2, wd // 2
    steps, jobs, chains, xops = [], [], [], {}
    macs = 0
    for ox0 in range(0, wo, g):
        pix = [ox0 + j for j in range(g) if ox0 + j < wo]
        chain_begin = len(jobs)
        init = [False] * g                 # this job's accumulators already written by the previous job's NEXT ops
        for oy in range(ho):
            sb = len(steps)
            init_next = [False] * g
            kys = [ky for ky in range(4) if (oy == 0 or ky >= 2)]      # rows with ky < 2 came from job oy - 1
            for ky in kys:
                iy = 2 * oy - 1 + ky
                if not 0 <= iy < h:
                    continue
                dual = ky >= 2 and oy + 1 < ho
                for ix in range(2 * pix[0] - 1, 2 * pix[-1] + 3):
                    if not 0 <= ix < wd:
                        continue
                    for kc in range(nkc):
                        ops = []
                        for j, ox in enumerate(pix):
                            kx = ix - (2 * ox - 1)
                            if not 0 <= kx < 4:
                                continue
                            ops.append((tile_of[(ky, kx, kc)] * cpo, j * cpo, STEP_ACCUM if init[j] else 0))
                            init[j] = True
                            if dual:
                                ops.append((tile_of[(ky - 2, kx, kc)] * cpo, j * cpo, STEP_NEXT | (STEP_ACCUM if init_next[j] else 0)))
                                init_next[j] = True
                        assert ops
                        b_row, d_col, fl = ops[0]
                        if len(ops) > 1:
                            xops[len(steps)] = ops[1:]
                        steps.append(((iy * wd + ix) * cpi + kc * KCHUNK, b_row, d_col, fl | STEP_B_RESIDENT))
            assert len(steps) > sb and all(init[j] for j in range(len(pix)))
            lo, hi = _mask_words([ox0 + j < wo for j in range(g)])
            jobs.append((sb, len(steps) - sb, (oy * wo + ox0) * cpo, 0, lo, hi, 0, 0))
            init = init_next
        chains.append((chain_begin, len(jobs) - chain_begin))
    for oy in range(ho):
        for ox in range(wo):
            macs += cin * cout * sum(1 for ky in range(4) for kx in range(4)
                                     if 0 <= 2 * oy - 1 + ky < h and 0 <= 2 * ox - 1 + kx < wd)
    col_map = list(range(nt))
    bias = [bp[n % cpo] for n in range(nt)]
    return _finish(name, cpo, nt, cpo, act, h * wd * cpi, ho * wo * cpo, 0, tiles, True, steps, jobs, col_map,
                   [-1] * nt, bias, macs_exact=macs, xops=xops, chains=chains,
                   meta=dict(kind="conv_k4s2p1_chained", out_hw=(ho, wo), out_c=cout, out_cp=cpo))


# ----------------------------------------------------------------------------------------------
# (d) Conv2d(1 -> Cout, k=4, s=2, p=1) on the padded bf16 28x28 image  (first discriminator /
#     encoder conv, src/utils/util_dcgan.py:110).  Cin = 1 gives K = 16 per output pixel, far
#     below an MMA K-chunk, so four image rows (4 x 32 padded columns = 128 contiguous
#     elements) are the K dimension and a banded weight matrix selects each output pixel's
#     4x4 window.  Starting the window 8 columns later per group of 4 output pixels makes the
#     band position-independent: the layer's whole B operand is 2 tiles, resident in SMEM.
# ----------------------------------------------------------------------------------------------

def lower_conv_image_k4s2p1(name, w, b, act) -> LayerProgram:
    cout = w.shape[0]
    assert w.shape[1] == 1 and tuple(w.shape[2:]) == (4, 4)
    cpo = cpitch(cout)
    gox = MAX_N // cpo                 # output pixels per job
    assert gox == 4, "window shift must be a multiple of 8 elements (4 output pixels)"
    nt = gox * cpo
    ho = wo = 14
    tiles = []
    for kc in range(2):
        t = np.zeros((nt, KCHUNK), np.float32)
        for kk in range(KCHUNK):
            r, j = divmod(kc * KCHUNK + kk, IMG_PAD_PITCH)      # window row 0..3, window column 0..31
            for oxl in range(gox):
                kx = j - 2 * oxl
                if 0 <= kx < 4:
                    t[oxl * cpo:oxl * cpo + cout, kk] = w[:, 0, r, kx]
        tiles.append(t)
    bp = np.zeros(cpo, np.float32)
    bp[:cout] = b
    # Bias through the MMA (see lower_convt_from_1x1): a third K-chunk per job reads padded rows 29-30, whose last two elements
    # (row 30, columns 30 and 31: no window multiplies them with a non-zero weight) hold 1.0 -- written by
    # DiscriminatorNet / EncoderNet.forward -- against the bias as a bf16 hi + lo pair.  The layer is bound by its epilogue's
    # instruction chain (two K-steps of MMA per 256 output columns), so the extra step is free and the epilogue loses its
    # bias loads and adds.
    ones_at = (IMG_PAD_ROWS - 1) * IMG_PAD_PITCH + IMG_PAD_PITCH - 2
    fold = BIAS_IN_K and BIAS_IN_K_IMAGE
    meta_extra = {}
    if fold:
        t = np.zeros((nt, KCHUNK), np.float32)
        hi_b = to_bf16_f32(bp)
        for oxl in range(gox):
            t[oxl * cpo:(oxl + 1) * cpo, KCHUNK - 2] = hi_b
            t[oxl * cpo:(oxl + 1) * cpo, KCHUNK - 1] = bp - hi_b
        tiles.append(t)
        bp = np.zeros(cpo, np.float32)
        meta_extra = dict(ones_cols=(ones_at, 2))
    steps, jobs = [], []
    for oy in range(ho):
        for g in range((wo + gox - 1) // gox):
            sb = len(steps)
            start = 2 * oy * IMG_PAD_PITCH + 2 * gox * g      # padded row 2oy == real row 2oy-1; padded col == real col-1+1
            for kc in range(2):
                steps.append((start + kc * KCHUNK, kc * nt, 0, STEP_B_RESIDENT | (STEP_ACCUM if kc else 0)))
            if fold:
                steps.append((ones_at + 2 - KCHUNK, 2 * nt, 0, STEP_B_RESIDENT | STEP_ACCUM))
            lo, hi = _mask_words([gox * g + j < wo for j in range(gox)])
            jobs.append((sb, len(steps) - sb, (oy * wo + gox * g) * cpo, 0, lo, hi, 0, 0))
    assert steps[-1][0] + KCHUNK <= IMG_PAD_ROWS * IMG_PAD_PITCH
    # exact MACs: taps inside the image
    taps = sum(1 for o in range(14) for kk in range(4) if 0 <= 2 * o - 1 + kk < 28)
    return _finish(name, nt, nt, cpo, act, IMG_PAD_ROWS * IMG_PAD_PITCH, ho * wo * cpo, 0, tiles, True, steps, jobs,
                   list(range(nt)), [-1] * nt, [bp[n % cpo] for n in range(nt)], macs_exact=taps * taps * cout,
                   meta=dict(kind="conv_image", out_hw=(ho, wo), out_c=cout, out_cp=cpo, **meta_extra))


# ----------------------------------------------------------------------------------------------
# (e) Conv2d(Cin -> Cout, k=7, s=2, p=0) on a 7x7 input == one dot product over the whole map
#     (discriminator score / encoder latent, src/utils/util_dcgan.py:113,
#      src/inverter/utils_ae/util_inverter.py:173).  N = Cout padded to 16; fp32 output.
# ----------------------------------------------------------------------------------------------

def lower_conv_full(name, w, b, act, in_cp=None, in_pitch=None) -> LayerProgram:
    cout, cin, k, _ = w.shape
    cpi = in_cp or cpitch(cin)
    nt = rup(cout, 16)
    assert nt <= MAX_N
    nkc = rup(cin, KCHUNK) // KCHUNK
    wp = np.zeros((nt, nkc * KCHUNK, k, k), np.float32)
    wp[:cout, :cin] = w
    tiles, steps = [], []
    for pix in range(k * k):
        ky, kx = divmod(pix, k)
        for kc in range(nkc):
            steps.append((pix * cpi + kc * KCHUNK, len(tiles) * nt, 0, STEP_ACCUM if steps else 0))
            tiles.append(wp[:, kc * KCHUNK:(kc + 1) * KCHUNK, ky, kx].copy())
    jobs = [(0, len(steps), 0, 0, 1, 0, 0, 0)]
    bias = np.zeros(nt, np.float32)
    bias[:cout] = b
    cm32 = [n if n < cout else -1 for n in range(nt)]
    return _finish(name, nt, nt, nt, act, max(k * k * cpi, in_pitch or 0), 0, cout, tiles, False,
                   steps, jobs, [-1] * nt, cm32, bias, macs_exact=cin * cout * k * k,
                   meta=dict(kind="conv_full", out_c=cout))


# ----------------------------------------------------------------------------------------------
# (f) General Conv2d(Cin -> Cout, k, stride, pad) on an NHWC activation -- the assessor's bottleneck convs
#     (src/pso/util_cnn.py:160-190: 1x1, 3x3 stride 1/2 pad 1, 1x1 stride-2 shortcut).  M = particles, so a
#     2x2 or 1x1 feature map costs no padding; border taps are simply absent.  Cout >= 256: one job per
#     (output pixel, 256-channel group), one N = 256 op per (tap, K-chunk).  Cout < 256: 256 / Cout consecutive
#     output pixels per job, each with its own N = Cout ops.  `residual`: the epilogue adds the bf16 tensor
#     res[P, same layout as the output] before the activation (x += identity; relu, :185-187).
# ----------------------------------------------------------------------------------------------

def lower_conv2d(name, w, b, act, h, wd, stride, pad, in_cp=None, residual=False, in_base=0, out_base=0, out_pitch=None,
                 shortcut=None, in_pitch=None) -> LayerProgram:
    """`in_base` / `out_base`: element offsets of the input map / the output map inside their particle rows (several
    maps may share one row-major buffer); `out_pitch`: row pitch of the output buffer (default: the map itself).
    `shortcut` folds the bottleneck's skip connection into the SAME accumulator as further MMA steps, so that its
    operand arrives through the pipelined TMA ring instead of the epilogue:
      ("identity", x_base)                      + x[pixel, channel]  (one 64x64 identity block per 64 channels, exact in fp32)
      ("conv", w_s, b_s, stride_s, x_base, h_x, wd_x, x_cp)   + the 1x1 projection conv of x (identity_downsample)."""
    cout, cin, k, k2 = w.shape
    assert k == k2
    cpi = in_cp or cpitch(cin)
    cpo = cpitch(cout)
    ho, wo = (h + 2 * pad - k) // stride + 1, (wd + 2 * pad - k) // stride + 1
    mma_n = cpo if cpo <= MAX_N else max(n for n in (256, 192, 128, 64) if cpo % n == 0)
    ncg = cpo // mma_n                  # channel groups per output pixel (Cout > 256)
    g = MAX_N // mma_n                  # output pixels per job (Cout < 256)
    nt = g * mma_n
    nkc = rup(cin, KCHUNK) // KCHUNK
    wp = np.zeros((cpo, nkc * KCHUNK, k, k), np.float32)
    wp[:cout, :cin] = w
    bp = np.zeros(cpo, np.float32)
    bp[:cout] = b
    in_need = max(in_base + h * wd * cpi, in_pitch or 0)      # in_pitch: row pitch of the (shared) input buffer
    sc_kind = shortcut[0] if shortcut else None
    if sc_kind == "identity":
        x_base = shortcut[1]
        assert cpo % KCHUNK == 0      # cout < cpo is fine: x's padding channels are 0, so the identity adds nothing there
        in_need = max(in_need, x_base + ho * wo * cpo)
    elif sc_kind == "conv":
        _, w_s, b_s, stride_s, x_base, h_x, wd_x, x_cp = shortcut
        assert w_s.shape[0] == cout and tuple(w_s.shape[2:]) == (1, 1)
        assert ((h_x - 1) // stride_s + 1, (wd_x - 1) // stride_s + 1) == (ho, wo)
        cin_s = w_s.shape[1]
        nkc_s = rup(cin_s, KCHUNK) // KCHUNK
        ws = np.zeros((cpo, nkc_s * KCHUNK), np.float32)
        ws[:cout, :cin_s] = w_s[:, :, 0, 0]
        bp[:cout] += b_s
        in_need = max(in_need, x_base + h_x * wd_x * x_cp)
    tiles, tile_of = [], {}

    def tile(key, make):
        if key not in tile_of:
            tile_of[key] = len(tiles)
            tiles.append(make())
        return tile_of[key] * mma_n

    steps, jobs = [], []
    macs = 0
    npix = ho * wo
    # Small weight packs stay RESIDENT in shared memory (no per-step weight tile through L2, which is what bounds the
    # N = 64 / 128 layers), and one A tile then feeds every output pixel of the job it touches (up to 4 ops per step).
    wbytes = k * k * nkc * cpo * KCHUNK * 2
    if CONV2D_RESIDENT and shortcut is None and not residual and ncg == 1 and wbytes <= CONV2D_RESIDENT_MAX:
        xops = {}
        merge = MERGE_OPS and stride == 1
        if merge:
            # taps of one kernel row back to back, kx DESCENDING: input pixel ix feeds output pixels ix - kx + pad in ascending order
            for ky in range(k):
                for kc in range(nkc):
                    for kx in reversed(range(k)):
                        tile((ky, kx, kc, 0), lambda: wp[:, kc * KCHUNK:(kc + 1) * KCHUNK, ky, kx].copy())
        for pix0 in range(0, npix, g):
            sb = len(steps)
            outs = [divmod(pix0 + j, wo) for j in range(min(g, npix - pix0))]
            feeds = {}                      # input pixel -> [(j, ky, kx)]
            for j, (oy, ox) in enumerate(outs):
                for ky in range(k):
                    for kx in range(k):
                        iy, ix = stride * oy - pad + ky, stride * ox - pad + kx
                        if 0 <= iy < h and 0 <= ix < wd:
                            feeds.setdefault((iy, ix), []).append((j, ky, kx))
                            macs += cin * cout
            started = set()
            for (iy, ix) in sorted(feeds):
                for kc in range(nkc):
                    ops = []
                    for (j, ky, kx) in feeds[(iy, ix)]:
                        b_row = tile((ky, kx, kc, 0), lambda: wp[:, kc * KCHUNK:(kc + 1) * KCHUNK, ky, kx].copy())
                        ops.append((b_row, j * mma_n, (STEP_ACCUM if j in started else 0) | STEP_B_RESIDENT))
                        started.add(j)
                    if merge:        # neighbouring columns x neighbouring weight rows, same accumulate flag -> one wider op
                        merged = []
                        for (br, dc, fl) in ops:
                            if merged:
                                pb, pd, pf = merged[-1]
                                wdt = op_width(pf)
                                if (pf & 0xFF) == fl and br == pb + wdt * mma_n and dc == pd + wdt * mma_n and wdt < 4:
                                    merged[-1] = (pb, pd, (pf & 0xFF) | (wdt << STEP_WIDE_SHIFT))
                                    continue
                            merged.append((br, dc, fl))
                        ops = merged
                    for o0 in range(0, len(ops), 4):
                        grp = ops[o0:o0 + 4]
                        steps.append(((iy * wd + ix) * cpi + in_base + kc * KCHUNK, grp[0][0], grp[0][1], grp[0][2]))
                        if len(grp) > 1:
                            xops[len(steps) - 1] = grp[1:]
            assert len(started) == len(outs), "an output pixel without any tap"
            lo, hi = _mask_words([pix0 + j < npix for j in range(g)])
            jobs.append((sb, len(steps) - sb, out_base + pix0 * cpo, 0, lo, hi, 0, 0))
        out_pitch = out_pitch or npix * cpo
        assert out_pitch >= out_base + npix * cpo
        return _finish(name, mma_n, nt, mma_n, act, in_need, out_pitch, 0, tiles, True, steps, jobs, list(range(nt)),
                       [-1] * nt, [[bp[n % mma_n] for n in range(nt)]], macs_exact=macs, xops=xops,
                       meta=dict(kind="conv2d", out_hw=(ho, wo), out_c=cout, out_cp=cpo, residual=False, out_base=out_base,
                                 out_elems=npix * cpo, shortcut=None, hbm_bytes=2 * (h * wd * cpi + npix * cpo)))
    for pix0 in range(0, npix, g):
        for cg in range(ncg):
            sb = len(steps)
            for j in range(min(g, npix - pix0)):
                oy, ox = divmod(pix0 + j, wo)
                first = True
                for ky in range(k):
                    iy = stride * oy - pad + ky
                    if not 0 <= iy < h:
                        continue
                    for kx in range(k):
                        ix = stride * ox - pad + kx
                        if not 0 <= ix < wd:
                            continue
                        if cg == 0:
                            macs += cin * cout
                        for kc in range(nkc):
                            b_row = tile((ky, kx, kc, cg), lambda: wp[cg * mma_n:(cg + 1) * mma_n, kc * KCHUNK:(kc + 1) * KCHUNK, ky, kx].copy())
                            steps.append((in_base + (iy * wd + ix) * cpi + kc * KCHUNK, b_row, j * mma_n, 0 if first else STEP_ACCUM))
                            first = False
                assert not first, "an output pixel without any tap"
                if sc_kind == "identity":
                    for i in range(mma_n // KCHUNK):      # channels [cg * mma_n + 64 i, + 64) of x feed columns [64 i, 64 i + 64)
                        a_col = x_base + (pix0 + j) * cpo + cg * mma_n + i * KCHUNK
                        if EYE_OPS:                       # N = 64 op against the engine's built-in identity tile: no weight traffic
                            steps.append((a_col, 0, j * mma_n + i * KCHUNK, STEP_ACCUM | STEP_EYE))
                            continue
                        def eye(i=i):
                            t = np.zeros((mma_n, KCHUNK), np.float32)
                            t[i * KCHUNK + np.arange(KCHUNK), np.arange(KCHUNK)] = 1.0
                            return t
                        steps.append((a_col, tile(("eye", i), eye), j * mma_n, STEP_ACCUM))
                elif sc_kind == "conv":
                    if cg == 0:
                        macs += cin_s * cout
                    for kc in range(nkc_s):
                        b_row = tile(("sc", kc, cg), lambda: ws[cg * mma_n:(cg + 1) * mma_n, kc * KCHUNK:(kc + 1) * KCHUNK].copy())
                        steps.append((x_base + ((stride_s * oy) * wd_x + stride_s * ox) * x_cp + kc * KCHUNK, b_row, j * mma_n, STEP_ACCUM))
            lo, hi = _mask_words([pix0 + j < npix for j in range(g)])
            jobs.append((sb, len(steps) - sb, out_base + pix0 * cpo + cg * mma_n, 0, lo, hi, cg, 0))
    bias = [[bp[cg * mma_n + (n % mma_n)] for n in range(nt)] for cg in range(ncg)]
    out_pitch = out_pitch or npix * cpo
    assert out_pitch >= out_base + npix * cpo
    return _finish(name, mma_n, nt, mma_n, act, in_need, out_pitch, 0, tiles, False, steps, jobs, list(range(nt)),
                   [-1] * nt, bias, macs_exact=macs,
                   meta=dict(kind="conv2d", out_hw=(ho, wo), out_c=cout, out_cp=cpo, residual=bool(residual), out_base=out_base,
                             out_elems=npix * cpo, shortcut=sc_kind,
                             # compulsory HBM traffic per particle: every input map once, the output once (bf16)
                             hbm_bytes=2 * (h * wd * cpi + npix * cpo + (npix * cpo if (sc_kind == "identity" or residual) else 0) +
                                            (shortcut[5] * shortcut[6] * shortcut[7] if sc_kind == "conv" else 0))))


# ----------------------------------------------------------------------------------------------
# (g) The assessor's stem: Conv2d(1 -> 64, k=7, s=2, p=3) on the 28x28 image (src/pso/util_cnn.py:85) as a
#     banded GEMM like (d): the image is stored zero-padded by 3 with a 40-element row pitch, a job owns 4
#     consecutive output pixels of one row and reads ONE contiguous window of 5 x 64 elements (8 padded rows
#     from (2 oy, 8 g)); the weight tiles are position-independent.
# ----------------------------------------------------------------------------------------------
STEM_PAD = 3
STEM_RESIDENT = knob("PSOB_STEM_RESIDENT", "0") != "0"     # 160 KB of stem weights resident over a CTA pair: measured no faster (MMA-bound on the zero-padded window)
STEM_PITCH = 40
STEM_ROWS = 36


def lower_conv_image_k7s2p3(name, w, b, act) -> LayerProgram:
    cout = w.shape[0]
    assert w.shape[1] == 1 and tuple(w.shape[2:]) == (7, 7)
    cpo = cpitch(cout)
    gox = MAX_N // cpo
    assert gox == 4, "window shift must be a multiple of 8 elements (4 output pixels)"
    nt = gox * cpo
    ho = wo = 14
    nk = 5
    tiles = []
    for kc in range(nk):
        t = np.zeros((nt, KCHUNK), np.float32)
        for kk in range(KCHUNK):
            r, j = divmod(kc * KCHUNK + kk, STEM_PITCH)
            if r >= 7:
                continue
            for oxl in range(gox):
                kx = j - 2 * oxl
                if 0 <= kx < 7:
                    t[oxl * cpo:oxl * cpo + cout, kk] = w[:, 0, r, kx]
        tiles.append(t)
    steps, jobs = [], []
    for oy in range(ho):
        for g in range((wo + gox - 1) // gox):
            sb = len(steps)
            start = 2 * oy * STEM_PITCH + 2 * gox * g
            for kc in range(nk):
                steps.append((start + kc * KCHUNK, kc * nt, 0, (STEP_ACCUM if kc else 0) | (STEP_B_RESIDENT if STEM_RESIDENT else 0)))
            lo, hi = _mask_words([gox * g + j < wo for j in range(gox)])
            jobs.append((sb, nk, (oy * wo + gox * g) * cpo, 0, lo, hi, 0, 0))
    assert steps[-1][0] + KCHUNK <= STEM_ROWS * STEM_PITCH
    bp = np.zeros(cpo, np.float32)
    bp[:cout] = b
    taps = sum(1 for o in range(14) for kk in range(7) if 0 <= 2 * o - 3 + kk < 28)
    return _finish(name, nt, nt, cpo, act, STEM_ROWS * STEM_PITCH, ho * wo * cpo, 0, tiles, STEM_RESIDENT, steps, jobs,
                   list(range(nt)), [-1] * nt, [bp[n % cpo] for n in range(nt)], macs_exact=taps * taps * cout,
                   meta=dict(kind="conv_image_k7", out_hw=(ho, wo), out_c=cout, out_cp=cpo,
                             hbm_bytes=2 * (STEM_ROWS * STEM_PITCH + ho * wo * cpo)))


def pad_image_stem(img: np.ndarray) -> np.ndarray:
    """[P, 28, 28] -> the stem's padded input rows [P, 36 * 40] (host-side twin of psob_pack_image_stem)."""
    p = img.shape[0]
    out = np.zeros((p, STEM_ROWS, STEM_PITCH), np.float32)
    out[:, STEM_PAD:STEM_PAD + 28, STEM_PAD:STEM_PAD + 28] = img.reshape(p, 28, 28)
    return out.reshape(p, -1)


def maxpool_nhwc(x: np.ndarray, h, wd, cp, k, stride, pad) -> np.ndarray:
    """MaxPool2d(k, stride, pad) on [P, h * wd * cp] (host-side twin of psob_maxpool_nhwc; padding never wins)."""
    p = x.shape[0]
    ho, wo = (h + 2 * pad - k) // stride + 1, (wd + 2 * pad - k) // stride + 1
    xi = x.reshape(p, h, wd, cp)
    out = np.full((p, ho, wo, cp), -np.inf, np.float32)
    for oy in range(ho):
        for ox in range(wo):
            y0, y1 = max(0, stride * oy - pad), min(h, stride * oy - pad + k)
            x0, x1 = max(0, stride * ox - pad), min(wd, stride * ox - pad + k)
            out[:, oy, ox] = xi[:, y0:y1, x0:x1].max(axis=(1, 2))
    return out.reshape(p, -1)


# ----------------------------------------------------------------------------------------------
# CPU emulation of the engine (numpy): the semantics csrc/pixel_gemm.cuh must reproduce.
# ----------------------------------------------------------------------------------------------

def _act(x, act):
    if act == ACT_RELU:
        return np.maximum(x, 0)
    if act == ACT_LEAKY:
        return np.where(x > 0, x, np.float32(0.2) * x)
    if act == ACT_TANH:
        return np.tanh(x)
    if act == ACT_SIGMOID:
        return 1.0 / (1.0 + np.exp(-x))
    return x


def emulate(prog: LayerProgram, x: np.ndarray, round_out=True, res: Optional[np.ndarray] = None):
    """x: float32 [P, in_pitch] holding bf16-representable values; res: [P, out_pitch] added before the
    activation (layers lowered with residual=True).  Returns (y_bf16_as_f32 or None, y_f32 or None)."""
    assert (res is not None) == bool(prog.meta.get("residual")), "residual operand does not match the layer"
    p = x.shape[0]
    assert x.shape[1] >= prog.in_pitch
    y = np.zeros((p, prog.out_pitch), np.float32) if prog.out_pitch else None
    y32 = np.zeros((p, prog.out32_pitch), np.float32) if prog.out32_pitch else None

    def apply(acc, a, b_row, d_col, flags):
        if flags & STEP_EYE:                 # identity op: 64 columns (+)= the A tile itself
            if flags & STEP_ACCUM:
                acc[:, d_col:d_col + KCHUNK] += a.astype(np.float32)
            else:
                acc[:, d_col:d_col + KCHUNK] = a.astype(np.float32)
            return
        n = prog.mma_n * op_width(flags)
        contrib = a.astype(np.float32) @ prog.wpack[b_row:b_row + n].T.astype(np.float32)
        if flags & STEP_ACCUM:
            acc[:, d_col:d_col + n] += contrib
        else:
            acc[:, d_col:d_col + n] = contrib

    for jb, jc in prog.chains:
        nxt = None
        for ji in range(jb, jb + jc):
            job = prog.jobs[ji]
            sb, sc, base, base32 = int(job[0]), int(job[1]), int(job[2]), int(job[3])
            mask = (int(np.uint32(job[4])) | (int(np.uint32(job[5])) << 32))
            acc = nxt if nxt is not None else np.zeros((p, prog.nt), np.float32)     # NEXT ops of the previous job landed here
            nxt = None
            for si in range(sb, sb + sc):
                a_col, b_row, d_col, flags = (int(v) for v in prog.steps[si])
                a = x[:, a_col:a_col + KCHUNK]
                for (br, dc, fl) in [(b_row, d_col, flags)] + [tuple(int(v) for v in o) for o in prog.xops.get(si, [])]:
                    if fl & STEP_NEXT:
                        assert ji + 1 < jb + jc, "NEXT op in the last job of a chain"
                        if nxt is None:
                            nxt = np.zeros((p, prog.nt), np.float32)
                        apply(nxt, a, br, dc, fl)
                    else:
                        apply(acc, a, br, dc, fl)
            bias = prog.bias[int(job[6])]
            pre = acc + bias[None, :]
            if res is not None:
                idx = base + prog.col_map.astype(np.int64)
                ok = (prog.col_map >= 0) & (idx < res.shape[1])
                pre[:, ok] = pre[:, ok] + res[:, idx[ok]]
            v = _act(pre, prog.act).astype(np.float32)
            n_cols = (mask >> 8) & 0xFF if prog.meta.get("band_width") else 0
            for n in range(prog.nt):
                pix = n // prog.cc
                if pix < 64 and not (mask >> pix) & 1:
                    continue
                if n_cols and not 1 <= n % prog.cc <= n_cols:      # tiled image band: columns past the image's right border
                    continue
                if y is not None and prog.col_map[n] >= 0:
                    y[:, base + prog.col_map[n]] = to_bf16_f32(v[:, n]) if round_out else v[:, n]
                if y32 is not None and prog.col_map32[n] >= 0:
                    y32[:, base32 + prog.col_map32[n]] = v[:, n]
    return y, y32
