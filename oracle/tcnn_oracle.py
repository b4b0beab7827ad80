"""CPU oracle for the HashGrid -> FullyFusedMLP -> L2 -> Adam hot path.  TEST INFRASTRUCTURE ONLY.

This file is a numpy restatement of the reference's algorithm (leejaeyong7/tiny-cuda-nn, a fork of
NVlabs/tiny-cuda-nn v1.7).  Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline /
--impl reference legs may import it, and only as the checker.  The product (tiny-cuda-nn_b200/) never
imports it and has no CPU fallback.

Parity pinning: the reference ships NO golden vectors / known-answer tests for this path,
and it has no CPU path, so it cannot run in the authoring container.  The oracle is pinned against dumps
of the reference's own CUDA build (oracle/_ref/ref_driver, built by oracle/build_ref.sh from the
unmodified sources) taken on a B200 and committed under tests/golden/ (see tests/golden/README.md and
tests/golden/make_golden.sh).  Until those fixtures exist the status is "parity unpinned".

Every function cites the reference file:line (relative to /root/reference) it restates.
"""
from __future__ import annotations

import math
from dataclasses import dataclass, field

import numpy as np

U32 = np.uint32
F32 = np.float32
F16 = np.float16

# include/tiny-cuda-nn/common_device.h:650-654 (coherent_prime_hash factors)
COHERENT_PRIMES = np.array([1, 2654435761, 805459861, 3674653429, 2097192037, 1434869437, 2165219737], dtype=np.uint64)

BATCH_SIZE_GRANULARITY = 256  # include/tiny-cuda-nn/common.h:235
LOSS_SCALE = 128.0            # include/tiny-cuda-nn/common.h:229-233 (default_loss_scale<__half>)


def next_multiple(v: int, m: int) -> int:
    return (v + m - 1) // m * m


# --------------------------------------------------------------------------------------------------
# pcg32 (dependencies/pcg32/pcg32.h:42-110,145-170: O'Neill's PCG-XSH-RR 64/32)
# --------------------------------------------------------------------------------------------------
PCG32_MULT = 0x5851F42D4C957F2D
M64 = (1 << 64) - 1


class Pcg32:
    def __init__(self, initstate: int = 0x853C49E6748FEA9B, initseq: int | None = None):
        if initseq is None and initstate == 0x853C49E6748FEA9B:
            # default ctor: fixed state/stream (pcg32.h:43)
            self.state = 0x853C49E6748FEA9B
            self.inc = 0xDA3E39CB94B95BDB
            return
        self.seed(initstate, 1 if initseq is None else initseq)

    def seed(self, initstate: int, initseq: int = 1):
        # pcg32.h:54-60
        self.state = 0
        self.inc = ((initseq << 1) | 1) & M64
        self.next_uint()
        self.state = (self.state + initstate) & M64
        self.next_uint()

    def next_uint(self) -> int:
        # pcg32.h:63-69
        old = self.state
        self.state = (old * PCG32_MULT + self.inc) & M64
        xorshifted = (((old >> 18) ^ old) >> 27) & 0xFFFFFFFF
        rot = old >> 59
        return ((xorshifted >> rot) | (xorshifted << ((-rot) & 31))) & 0xFFFFFFFF

    def advance(self, delta: int):
        # pcg32.h:145-168 (Brown's arbitrary-stride jump)
        cur_mult, cur_plus, acc_mult, acc_plus = PCG32_MULT, self.inc, 1, 0
        delta &= M64
        while delta > 0:
            if delta & 1:
                acc_mult = (acc_mult * cur_mult) & M64
                acc_plus = (acc_plus * cur_mult + cur_plus) & M64
            cur_plus = ((cur_mult + 1) * cur_plus) & M64
            cur_mult = (cur_mult * cur_mult) & M64
            delta >>= 1
        self.state = (acc_mult * self.state + acc_plus) & M64

    def copy(self) -> "Pcg32":
        r = Pcg32.__new__(Pcg32)
        r.state, r.inc = self.state, self.inc
        return r


def _pcg_output(old: np.ndarray) -> np.ndarray:
    xorshifted = (((old >> np.uint64(18)) ^ old) >> np.uint64(27)).astype(np.uint32)
    rot = (old >> np.uint64(59)).astype(np.uint32)
    return (xorshifted >> rot) | (xorshifted << ((~rot + U32(1)) & U32(31)))


def _uint_to_float01(u: np.ndarray) -> np.ndarray:
    # pcg32.h:101-110 next_float: [1,2) mantissa trick
    return ((u >> U32(9)) | U32(0x3F800000)).view(np.float32) - F32(1.0)


def pcg32_stream(rng: Pcg32, n: int) -> np.ndarray:
    """n consecutive next_uint() outputs, vectorised (LCG states via repeated doubling); advances rng."""
    if n == 0:
        return np.zeros(0, np.uint32)
    with np.errstate(over="ignore"):
        states = np.empty(n, dtype=np.uint64)
        states[0] = np.uint64(rng.state)
        mult, plus, filled = np.uint64(PCG32_MULT), np.uint64(rng.inc), 1
        while filled < n:
            m = min(filled, n - filled)
            states[filled:filled + m] = states[:m] * mult + plus
            plus = (mult + np.uint64(1)) * plus
            mult = mult * mult
            filled += m
        out = _pcg_output(states)
    rng.advance(n)
    return out


def seed_seq_first(seed: int) -> int:
    """First 32-bit word of std::seed_seq{seed}.generate(2 words) (trainer.h:52-55; ISO C++ [rand.util.seedseq])."""
    n, s = 2, 1
    v = [seed & 0xFFFFFFFF]
    b = [0x8B8B8B8B] * n
    t = 0  # n < 7
    p = (n - t) // 2
    q = p + t
    m = max(s + 1, n)
    M32 = 0xFFFFFFFF

    def T(x):
        return x ^ (x >> 27)

    for k in range(m):
        r1 = (1664525 * T(b[k % n] ^ b[(k + p) % n] ^ b[(k - 1) % n])) & M32
        if k == 0:
            r2 = r1 + s
        elif k <= s:
            r2 = r1 + (k % n) + v[k - 1]
        else:
            r2 = r1 + (k % n)
        r2 &= M32
        b[(k + p) % n] = (b[(k + p) % n] + r1) & M32
        b[(k + q) % n] = (b[(k + q) % n] + r2) & M32
        b[k % n] = r2
    for k in range(m, m + n):
        r3 = (1566083941 * T((b[k % n] + b[(k + p) % n] + b[(k - 1) % n]) & M32)) & M32
        r4 = (r3 - (k % n)) & M32
        b[(k + p) % n] ^= r3
        b[(k + q) % n] ^= r4
        b[k % n] = r4
    return b[0]


# --------------------------------------------------------------------------------------------------
# Grid encoding
# --------------------------------------------------------------------------------------------------
def grid_scale(level: int, log2_per_level_scale: float, base_resolution: int) -> np.float32:
    # common_device.h:709-714  exp2f(level * log2_per_level_scale) * base_resolution - 1.0f   (all fp32)
    e = F32(F32(level) * F32(log2_per_level_scale))
    return F32(F32(np.exp2(e, dtype=np.float32) * F32(base_resolution)) - F32(1.0))


def grid_resolution(scale: np.float32) -> int:
    # common_device.h:716-718
    return int(np.ceil(F32(scale))) + 1


@dataclass
class GridConfig:
    """encodings/grid.h:1143-1208 JSON keys and defaults (HashGrid, CoherentPrime hash, Linear interpolation)."""
    n_dims: int
    n_levels: int = 16
    n_features_per_level: int = 2
    log2_hashmap_size: int = 19
    base_resolution: int = 16
    per_level_scale: float = 2.0
    # optional per-level (scale, resolution) table dumped from the device (CUDA exp2f vs libm)
    scales: np.ndarray | None = None
    grid_type: str = "Hash"           # Hash | Dense | Tiled (grid.h:702-712; DenseGrid / TiledGrid otypes, grid.h:1180-1208)
    max_level: float = 1000.0         # GridEncoding::m_max_level (grid_interface.h:120)
    stochastic: bool = False          # "stochastic_interpolation" (grid.h:284-303): backward pass only
    interpolation: str = "Linear"     # Nearest | Linear | Smoothstep (common_device.h:801-807; Nearest: grid.h:121-139)
    hash: str = "CoherentPrime"       # Prime | CoherentPrime | ReversedPrime | Rng (common_device.h:631-688, grid.h:1151)
    max_level_gpu: np.ndarray | None = None   # per-sample max level (GridEncoding::set_max_level_gpu, grid.h:69-75): fp32 [B]
    offsets: np.ndarray = field(init=False)
    resolutions: np.ndarray = field(init=False)

    def __post_init__(self):
        # grid.h:692-716 offset table (host-side, std::log2(float) then the host grid_scale)
        log2s = F32(np.log2(F32(self.per_level_scale)))
        self.log2_per_level_scale = log2s
        if self.scales is None:
            self.scales = np.array([grid_scale(l, log2s, self.base_resolution) for l in range(self.n_levels)], dtype=np.float32)
        self.resolutions = np.array([grid_resolution(s) for s in self.scales], dtype=np.uint32)
        offs, off = [], 0
        max_params = 0xFFFFFFFF // 2
        for l in range(self.n_levels):
            res = int(self.resolutions[l])
            dense = max_params if float(F32(res) ** F32(self.n_dims)) > float(F32(max_params)) else (res ** self.n_dims) & 0xFFFFFFFF
            n = next_multiple(dense, 8)
            gt = self.grid_type.lower()
            if gt == "hash":
                n = min(n, 1 << self.log2_hashmap_size)
            elif gt == "tiled":
                n = min(n, self.base_resolution ** self.n_dims)
            else:
                assert gt == "dense", self.grid_type
            offs.append(off)
            off += n
        offs.append(off)
        self.offsets = np.array(offs, dtype=np.uint32)

    @property
    def n_features(self) -> int:
        return self.n_levels * self.n_features_per_level

    @property
    def n_params(self) -> int:
        return int(self.offsets[-1]) * self.n_features_per_level

    def level_size(self, l: int) -> int:
        return int(self.offsets[l + 1]) - int(self.offsets[l])

    @property
    def hashed(self) -> bool:
        return self.grid_type.lower() == "hash"

    @property
    def smooth(self) -> bool:
        return self.interpolation.lower() == "smoothstep"

    @property
    def active_levels(self) -> int:
        """GridEncoding::set_max_level (grid_interface.h:101-107): levels with level >= max_level * n_levels + 1e-3 (fp32,
        grid.h:69-75) encode to zero and receive no gradient (grid.h:236-243).  Default 1000 = all levels."""
        F = F32(self.n_features_per_level)
        scaled = F32(F32(F32(self.max_level) * F32(self.n_levels * self.n_features_per_level)) / F)
        return sum(1 for l in range(self.n_levels) if not (F32(l) >= F32(scaled + F32(1e-3))))

    @property
    def nearest(self) -> bool:
        return self.interpolation.lower() == "nearest"


def sample_level_mask(cfg: GridConfig, B: int, level: int, backward: bool = False) -> np.ndarray:
    """Which samples see `level` when a per-sample max level is set (grid.h:69-75 forward: zero if level >= m + 1e-3;
    grid.h:236-243 backward: skipped if level > m + 1e-3; m = max_level_gpu[i] * (L * F) / F in fp32)."""
    if cfg.max_level_gpu is None:
        return np.ones(B, dtype=bool)
    F = F32(cfg.n_features_per_level)
    m = ((cfg.max_level_gpu[:B].astype(np.float32) * F32(cfg.n_levels * cfg.n_features_per_level)).astype(np.float32) / F).astype(np.float32)
    lim = (m + F32(1e-3)).astype(np.float32)
    return ~(F32(level) > lim) if backward else ~(F32(level) >= lim)


def smoothstep(v: np.ndarray) -> np.ndarray:
    """common_device.h:801-803, fp32."""
    v = v.astype(np.float32)
    return (v * v * (F32(3.0) - F32(2.0) * v)).astype(np.float32)


def smoothstep_derivative(v: np.ndarray) -> np.ndarray:
    """common_device.h:805-807."""
    v = v.astype(np.float32)
    return (F32(6.0) * v * (F32(1.0) - v)).astype(np.float32)


def pos_fract(x: np.ndarray, scale: np.float32):
    """common_device.h:857-868: pos = fmaf(scale, x, 0.5f); cell = (uint32)(int)floorf(pos); frac = pos - floorf(pos)."""
    # fmaf = single rounding of the exact scale*x+0.5; float64 holds the exact product of two fp32 and the sum's
    # rounding to fp32 from float64 is a double rounding only in measure-zero ties we accept here.
    p = (np.float64(scale) * x.astype(np.float64) + 0.5).astype(np.float32)
    fl = np.floor(p)
    cell = (fl.astype(np.int64) & 0xFFFFFFFF).astype(np.uint32)  # (uint32_t)(int)floorf: two's complement wrap
    frac = (p - fl).astype(np.float32)
    return frac, cell


HASH_FACTORS = {  # common_device.h:643-659
    "prime": [1958374283, 2654435761, 805459861, 3674653429, 2097192037, 1434869437, 2165219737],
    "coherentprime": [1, 2654435761, 805459861, 3674653429, 2097192037, 1434869437, 2165219737],
    "reversedprime": [2165219737, 1434869437, 2097192037, 3674653429, 805459861, 2654435761, 1958374283],
}


def grid_hash(pos_grid: np.ndarray, hash_type: str = "CoherentPrime") -> np.ndarray:
    """grid_hash, common_device.h:631-688: xor of pos[d] * factor[d] (uint32 wrap) for the three prime hashes; Rng: pcg32{1337} advanced
    by xor_d(pos[d] << d * (64 / D)) (a 64-bit step), next_uint()."""
    D = pos_grid.shape[-1]
    name = hash_type.lower()
    if name == "rng":
        bits = 64 // D
        step = np.zeros(pos_grid.shape[:-1], dtype=np.uint64)
        for d in range(D):
            step ^= (pos_grid[..., d].astype(np.uint64) << np.uint64(d * bits)) if d * bits < 64 else np.uint64(0)
        flat = step.reshape(-1)
        out = np.empty(flat.shape, dtype=np.uint32)
        for k, st in enumerate(flat):
            r = Pcg32(1337)
            r.advance(int(st) - (1 << 64) if int(st) >= (1 << 63) else int(st))  # (int64_t)step
            out[k] = r.next_uint()
        return out.reshape(step.shape).astype(np.uint64)
    f = HASH_FACTORS[name]
    h = np.zeros(pos_grid.shape[:-1], dtype=np.uint64)
    for d in range(D):
        h ^= (pos_grid[..., d].astype(np.uint64) * np.uint64(f[d])) & np.uint64(0xFFFFFFFF)
    return h


def grid_index(pos_grid: np.ndarray, hashmap_size: int, resolution: int, hashed: bool = True, hash_type: str = "CoherentPrime") -> np.ndarray:
    """common_device.h:690-707 for GridType::Hash, including the uint32 stride wrap.

    pos_grid: uint32 [..., D] -> uint32 [...]."""
    D = pos_grid.shape[-1]
    stride = 1
    index = np.zeros(pos_grid.shape[:-1], dtype=np.uint64)
    for d in range(D):
        if not (stride <= hashmap_size):
            break
        index = (index + pos_grid[..., d].astype(np.uint64) * np.uint64(stride)) & np.uint64(0xFFFFFFFF)
        stride = (stride * resolution) & 0xFFFFFFFF
    if hashed and hashmap_size < stride:
        index = grid_hash(pos_grid, hash_type)
    return (index % np.uint64(hashmap_size)).astype(np.uint32)


def grid_corners(cfg: GridConfig, x: np.ndarray, level: int):
    """Per-level corner indices (entry units, relative to the level) and fp32 weights.

    grid.h:94-163: corner idx bit d selects +1 in dim d; weight = prod(dim) of (1-frac) or frac, in fp32,
    multiplied in dim order starting from 1.0f.   x: fp32 [B, D] -> (uint32 [B, 2^D], fp32 [B, 2^D])."""
    D = cfg.n_dims
    frac, cell = pos_fract(x, cfg.scales[level])
    if cfg.nearest:  # grid.h:121-139 / 279-282: the entry at pos_grid alone, weight 1 (no interpolation)
        hs, res = cfg.level_size(level), int(cfg.resolutions[level])
        return grid_index(cell, hs, res, cfg.hashed, cfg.hash)[:, None], np.ones((x.shape[0], 1), dtype=np.float32)
    if cfg.smooth:
        frac = smoothstep(frac)
    B = x.shape[0]
    idx = np.empty((B, 1 << D), dtype=np.uint32)
    w = np.empty((B, 1 << D), dtype=np.float32)
    hs, res = cfg.level_size(level), int(cfg.resolutions[level])
    for c in range(1 << D):
        weight = np.ones(B, dtype=np.float32)
        pg = np.empty((B, D), dtype=np.uint32)
        for d in range(D):
            if (c >> d) & 1:
                weight = (weight * frac[:, d]).astype(np.float32)
                pg[:, d] = cell[:, d] + U32(1)
            else:
                weight = (weight * (F32(1) - frac[:, d])).astype(np.float32)
                pg[:, d] = cell[:, d]
        idx[:, c] = grid_index(pg, hs, res, cfg.hashed, cfg.hash)
        w[:, c] = weight
    return idx, w


def grid_forward(cfg: GridConfig, x: np.ndarray, table: np.ndarray) -> np.ndarray:
    """kernel_grid, grid.h:48-169.  table: fp16 [n_params]; returns fp16 [B, L*F] (sample-major).

    Interpolation in fp16: result = __hfma2((half)weight, value, result) per corner (grid.h:146-163, vec.h:370-375):
    one rounding per fma, emulated as float64 fma -> fp16."""
    B, F = x.shape[0], cfg.n_features_per_level
    out = np.zeros((B, cfg.n_features), dtype=np.float16)
    t = table.astype(np.float16).reshape(-1, F)
    for l in range(cfg.active_levels):
        idx, w = grid_corners(cfg, x, l)
        base = int(cfg.offsets[l])
        keep = sample_level_mask(cfg, B, l)
        if cfg.nearest:  # the entry itself, no arithmetic (grid.h:121-129) - keeps -0.0
            out[:, l * F:(l + 1) * F] = np.where(keep[:, None], t[base + idx[:, 0].astype(np.int64)], np.float16(0))
            continue
        res = np.zeros((B, F), dtype=np.float16)
        for c in range(idx.shape[1]):
            val = t[base + idx[:, c].astype(np.int64)]
            wh = w[:, c].astype(np.float16)
            res = (wh.astype(np.float64)[:, None] * val.astype(np.float64) + res.astype(np.float64)).astype(np.float16)
        out[:, l * F:(l + 1) * F] = np.where(keep[:, None], res, np.float16(0))
    return out


def grid_backward(cfg: GridConfig, x: np.ndarray, dy: np.ndarray, sequential_fp16: bool = False) -> np.ndarray:
    """kernel_grid_backward, grid.h:214-320.  dy: fp16 [B, L*F] -> float64 [n_params] (exact sum of the fp16
    products (half)weight * grad; the reference accumulates them with fp16 red atomics in arbitrary order, so
    comparisons are tolerance-based)."""
    F = cfg.n_features_per_level
    g = np.zeros((int(cfg.offsets[-1]), F), dtype=np.float64)
    for l in range(cfg.active_levels):
        keep = sample_level_mask(cfg, x.shape[0], l, backward=True)
        if cfg.stochastic and not cfg.nearest:
            # grid.h:284-303: sample = random_val(1337, i + level * B) = pcg32{1337} advanced by that (uint32) index; the same
            # draw picks cell or cell + 1 in every dimension; the whole gradient goes to that corner (weight 1)
            B = x.shape[0]
            frac, cell = pos_fract(x, cfg.scales[l])
            if cfg.smooth:
                frac = smoothstep(frac)
            sample = np.empty(B, np.float32)
            for i in range(B):
                r = Pcg32(1337)
                r.advance((i + l * B) & 0xFFFFFFFF)
                sample[i] = _uint_to_float01(np.array([r.next_uint()], dtype=np.uint32))[0]
            pg = np.where(sample[:, None] >= frac, cell, cell + U32(1)).astype(np.uint32)
            ids = grid_index(pg, cfg.level_size(l), int(cfg.resolutions[l]), cfg.hashed, cfg.hash).astype(np.int64)
            np.add.at(g, int(cfg.offsets[l]) + ids[keep], dy[keep, l * F:(l + 1) * F].astype(np.float16).astype(np.float64))
            continue
        idx, w = grid_corners(cfg, x, l)
        base = int(cfg.offsets[l])
        gl = dy[:, l * F:(l + 1) * F].astype(np.float16)
        contrib = []
        for c in range(idx.shape[1]):
            wh = w[:, c].astype(np.float16)
            prod = (wh.astype(np.float32)[:, None] * gl.astype(np.float32)).astype(np.float16)  # __hmul2, one rounding
            if F == 1:  # grad_t = float (grid.h:655-666): (float)weight * grad in fp32, fp32 atomics, one cast to fp16 at the end (grid.h:877-885)
                prod = (w[:, c].astype(np.float32)[:, None] * gl.astype(np.float32)).astype(np.float32)
            if not sequential_fp16:
                np.add.at(g, base + idx[keep, c].astype(np.int64), prod[keep].astype(np.float64))
            else:
                contrib.append((idx[:, c].astype(np.int64), prod))
        if sequential_fp16:
            # what red.add.noftz.f16x2 does: the running sum is rounded to fp16 after EVERY addition (order: sample, corner)
            rows = np.concatenate([cc[0] for cc in contrib]); vals = np.concatenate([cc[1] for cc in contrib])
            order = np.argsort(rows, kind="stable"); rows, vals = rows[order], vals[order]
            first = np.r_[True, rows[1:] != rows[:-1]]
            start = np.flatnonzero(first); rank = np.arange(len(rows)) - np.repeat(start, np.diff(np.r_[start, len(rows)]))
            acc = np.zeros((cfg.level_size(l), F), np.float16)
            for r in range(int(rank.max()) + 1):
                sel = rank == r
                acc[rows[sel]] = (acc[rows[sel]].astype(np.float32) + vals[sel].astype(np.float32)).astype(np.float16)
            g[base:base + cfg.level_size(l)] = acc.astype(np.float64)
    return g.reshape(-1)


def grid_forward_f32(cfg: GridConfig, x: np.ndarray, table: np.ndarray) -> np.ndarray:
    """kernel_grid with T = float (an encoding created with Precision::Fp32, cpp_api.cu:156-165): fp32 table, result =
    fmaf(weight, value, result) per corner in corner order (grid.h:146-163).  Returns fp32 [B, L*F]."""
    B, F = x.shape[0], cfg.n_features_per_level
    out = np.zeros((B, cfg.n_features), dtype=np.float32)
    t = table.astype(np.float32).reshape(-1, F)
    for l in range(cfg.active_levels):
        idx, w = grid_corners(cfg, x, l)
        base = int(cfg.offsets[l])
        if cfg.nearest:
            out[:, l * F:(l + 1) * F] = t[base + idx[:, 0].astype(np.int64)]
            continue
        res = np.zeros((B, F), dtype=np.float32)
        for c in range(idx.shape[1]):
            val = t[base + idx[:, c].astype(np.int64)]
            res = (w[:, c].astype(np.float64)[:, None] * val.astype(np.float64) + res.astype(np.float64)).astype(np.float32)
        out[:, l * F:(l + 1) * F] = res
    return out


def grid_backward_f32(cfg: GridConfig, x: np.ndarray, dy: np.ndarray) -> np.ndarray:
    """kernel_grid_backward with T = float: atomicAdd of the fp32 products weight * grad (grid.h:248-252), summed here in
    float64 (the atomics' order is arbitrary).  dy fp32 [B, >= L*F] -> float64 [n_params]."""
    F = cfg.n_features_per_level
    g = np.zeros((int(cfg.offsets[-1]), F), dtype=np.float64)
    for l in range(cfg.active_levels):
        idx, w = grid_corners(cfg, x, l)
        base = int(cfg.offsets[l])
        gl = dy[:, l * F:(l + 1) * F].astype(np.float32)
        for c in range(idx.shape[1]):
            prod = (w[:, c].astype(np.float32)[:, None] * gl).astype(np.float32)
            np.add.at(g, base + idx[:, c].astype(np.int64), prod.astype(np.float64))
    return g.reshape(-1)


def grid_backward_input(cfg: GridConfig, x: np.ndarray, table: np.ndarray, dy: np.ndarray, fp32: bool = False) -> np.ndarray:
    """dL/dx of the grid encoding: dy/dx as kernel_grid computes it when asked (grid.h:171-211: per level and feature,
    sum over the 2^(D-1) corner pairs along the gradient dim of scale * prod(other weights) * (val_right - val_left), fp32)
    contracted with dL/dy as kernel_grid_backward_input does (grid.h:322-349: k ascending).  x fp32 [B, D], table fp16,
    dy fp16 [B, >= L*F] -> fp32 [B, D]."""
    B, D = x.shape
    F = cfg.n_features_per_level
    t = table.astype(np.float32 if fp32 else np.float16).reshape(-1, F)  # fp32: create_encoding<float> (cpp_api.cu:156-165)
    result = np.zeros((B, D), np.float32)
    if cfg.nearest:  # "Gradient is zero when there's no interpolation" (grid.h:131-137)
        return result
    for l in range(cfg.active_levels):
        scale, res = F32(cfg.scales[l]), int(cfg.resolutions[l])
        frac, cell = pos_fract(x, F32(scale))
        dfrac = smoothstep_derivative(frac) if cfg.smooth else np.ones_like(frac)  # pos_derivative (grid.h:100-113)
        if cfg.smooth:
            frac = smoothstep(frac)
        size, base = cfg.level_size(l), int(cfg.offsets[l])
        grads = np.zeros((B, F, D), np.float32)
        for gd in range(D):
            for idx in range(1 << (D - 1)):
                weight = np.full(B, F32(scale), np.float32)
                pg = np.empty((B, D), np.uint32)
                for nd in range(D - 1):
                    dim = nd + 1 if nd >= gd else nd
                    if (idx & (1 << nd)) == 0:
                        weight = (weight * (F32(1.0) - frac[:, dim])).astype(np.float32); pg[:, dim] = cell[:, dim]
                    else:
                        weight = (weight * frac[:, dim]).astype(np.float32); pg[:, dim] = cell[:, dim] + np.uint32(1)
                pg[:, gd] = cell[:, gd]
                vl = t[base + grid_index(pg, size, res, cfg.hashed, cfg.hash).astype(np.int64)].astype(np.float32)
                pg[:, gd] = cell[:, gd] + np.uint32(1)
                vr = t[base + grid_index(pg, size, res, cfg.hashed, cfg.hash).astype(np.int64)].astype(np.float32)
                grads[:, :, gd] = (grads[:, :, gd] + (weight[:, None] * (vr - vl)).astype(np.float32) * dfrac[:, gd, None]).astype(np.float32)
        grads[~sample_level_mask(cfg, B, l)] = 0  # dy/dx of a level the sample does not see is zero (grid.h:78-89)
        for f in range(F):
            dl = dy[:, l * F + f].astype(np.float32)
            result = (result + dl[:, None] * grads[:, f, :]).astype(np.float32)
    return result


# --------------------------------------------------------------------------------------------------
# FullyFusedMLP (src/fully_fused_mlp.cu:635-678 weight layout, :499-557 forward, :150-259/:735-836 backward)
# --------------------------------------------------------------------------------------------------
@dataclass
class MlpConfig:
    """src/network.cu:107-113 defaults; weights are row-major [out, in], concatenated:
    [W x in_pad], (H-1) x [W x W], [out_pad16 x W]  (fully_fused_mlp.cu:654-671)."""
    n_input: int            # padded encoding width (multiple of 16)
    n_output: int
    n_neurons: int = 64
    n_hidden_layers: int = 2
    activation: str = "ReLU"
    output_activation: str = "None"

    @property
    def padded_output(self) -> int:
        return next_multiple(self.n_output, 16)

    def shapes(self):
        W = self.n_neurons
        return [(W, self.n_input)] + [(W, W)] * (self.n_hidden_layers - 1) + [(self.padded_output, W)]

    @property
    def n_params(self) -> int:
        return sum(a * b for a, b in self.shapes())

    def split(self, params: np.ndarray):
        out, pos = [], 0
        for a, b in self.shapes():
            out.append(params[pos:pos + a * b].reshape(a, b))
            pos += a * b
        return out


K_ACT = F32(10.0)  # common_device.h:100


def activation_forward(z16: np.ndarray, name: str) -> np.ndarray:
    """warp_activation<__half> (common_device.h:102-160): fp16 pre-activation in, function in fp32, fp16 out."""
    n = name.lower()
    x = z16.astype(np.float16).astype(np.float32)
    if n == "none":
        return z16.astype(np.float16)
    if n == "relu":
        return np.maximum(z16.astype(np.float16), np.float16(0))
    if n == "leakyrelu":  # half multiply by (half)1 or (half)0.01
        return (x * np.where(x > 0, np.float16(1.0), np.float16(0.01)).astype(np.float32)).astype(np.float16)
    with np.errstate(over="ignore"):
        if n == "exponential":
            r = np.exp(x, dtype=np.float32)
        elif n == "sigmoid":
            r = F32(1.0) / (F32(1.0) + np.exp(-x, dtype=np.float32))
        elif n == "squareplus":
            t = x * K_ACT
            r = F32(0.5) * (t + np.sqrt(t * t + F32(4.0), dtype=np.float32)) / K_ACT
        elif n == "softplus":
            r = np.log(np.exp(x * K_ACT, dtype=np.float32) + F32(1.0), dtype=np.float32) / K_ACT
        elif n == "tanh":
            r = np.tanh(x, dtype=np.float32)
        else:
            raise ValueError(name)
    return r.astype(np.float16)


def activation_backward(g16: np.ndarray, y16: np.ndarray, name: str) -> np.ndarray:
    """warp_activation_backward<__half> (common_device.h:240-293): fp16 gradient times an fp16 derivative factor that is
    computed from the stored POST-activation y."""
    n = name.lower()
    y = y16.astype(np.float16).astype(np.float32)
    if n == "none":
        return g16.astype(np.float16)
    if n == "relu":
        d = (y > 0).astype(np.float16)
    elif n == "leakyrelu":
        d = np.where(y > 0, np.float16(1.0), np.float16(0.01)).astype(np.float16)
    elif n == "exponential":
        d = y16.astype(np.float16)
    elif n == "sigmoid":  # (T)(y * (T)(1.0f - (float)y)): half * half
        d = (y * (F32(1.0) - y).astype(np.float16).astype(np.float32)).astype(np.float16)
    elif n == "squareplus":
        t = y * K_ACT
        d = (t * t / (t * t + F32(1.0))).astype(np.float16)
    elif n == "softplus":
        d = (F32(1.0) - np.exp(-y * K_ACT, dtype=np.float32)).astype(np.float16)
    elif n == "tanh":
        d = (F32(1.0) - y * y).astype(np.float16)
    else:
        raise ValueError(name)
    return (g16.astype(np.float16).astype(np.float32) * d.astype(np.float32)).astype(np.float16)


def matmul_acc(a: np.ndarray, b: np.ndarray, acc=np.float32) -> np.ndarray:
    """a [M, K] @ b [K, N] for fp16 operands.  acc = np.float32 / np.float64: one accumulation, the caller rounds.
    acc = "hmma16": the accumulator is rounded to fp16 after every 16-deep MMA step, which is how tensor-core fp16
    accumulation (the reference's wmma fragments, fully_fused_mlp.cu:47-129) behaves; numpy's own fp16 matmul does NOT do
    this (it accumulates in fp32 and rounds once)."""
    if acc != "hmma16":
        return a.astype(acc) @ b.astype(acc)
    out = np.zeros((a.shape[0], b.shape[1]), np.float16)
    for k0 in range(0, a.shape[1], 16):
        out = (out.astype(np.float32) + a[:, k0:k0 + 16].astype(np.float32) @ b[k0:k0 + 16].astype(np.float32)).astype(np.float16)
    return out


def _dw_splitk16(g: np.ndarray, inp: np.ndarray, slice_samples: int = 4096) -> np.ndarray:
    """g^T [out, B] @ inp [B, in] the way a split-K GEMM with fp16 accumulators behaves: every slice of `slice_samples` samples is
    accumulated in fp16 (rounded after each 16-sample step), the slices are then summed (fully_fused_mlp.cu:775: split_k = B / 4096)."""
    total = np.zeros((g.shape[1], inp.shape[1]), np.float64)
    for s0 in range(0, g.shape[0], slice_samples):
        total += matmul_acc(np.ascontiguousarray(g[s0:s0 + slice_samples].T), inp[s0:s0 + slice_samples], "hmma16").astype(np.float64)
    return total.astype(np.float16).astype(np.float64)


def mlp_forward(cfg: MlpConfig, params: np.ndarray, x: np.ndarray, acc="hmma16"):
    """x: fp16 [B, n_input] -> (output fp16 [B, out_pad], list of post-activation hidden fp16 [B, W]).

    Storage fp16 after every layer like the reference (fully_fused_mlp.cu:47-129); `acc` is the accumulator type.  The
    default "hmma16" is the reference's: fp16 wmma fragments, rounded after every 16-deep step (this reproduces the reference
    build's network output to <= 3 last-place differences per 3000 values; the B200 kernels accumulate the same way in TMEM
    and match it bit for bit).  np.float32 / np.float64 give the exactly-accumulated result for comparison."""
    ws = cfg.split(params.astype(np.float16))
    h = x.astype(np.float16)
    hidden = []
    for w in ws[:-1]:
        z = matmul_acc(h, w.T, acc)
        h = activation_forward(z.astype(np.float16), cfg.activation)
        hidden.append(h)
    y = activation_forward(matmul_acc(h, ws[-1].T, acc).astype(np.float16), cfg.output_activation)
    return y, hidden


def mlp_backward(cfg: MlpConfig, params: np.ndarray, x: np.ndarray, hidden, dy: np.ndarray, acc="hmma16", dw_acc=np.float64, y=None):
    """dy: fp16 [B, out_pad] (loss-scaled) -> (dparams float64 [n_params] exact-ish, dx fp16 [B, n_input]).

    Chain (fully_fused_mlp.cu:150-259): g = (W_next^T dy) * relu'(h) with the mask taken from the stored
    post-activation (common_device.h:243-248), rounded to fp16 at every layer (accumulator `acc` as in mlp_forward); weight
    gradients dW = g^T h_prev over the batch (fully_fused_mlp.cu:783-829).  The reference accumulates those in fp16 too, in
    split-K slices of 4096 samples (cutlass_matmul.h:437-479) whose summation order is CUTLASS's; `dw_acc` = np.float64 returns
    the exact sums, "splitk16" an emulation of the sliced fp16 accumulation (16-sample steps) that shows its underflow."""
    ws = cfg.split(params.astype(np.float16))
    grads = [None] * len(ws)
    g = dy.astype(np.float16)
    if cfg.output_activation.lower() != "none":
        # activation_backward_output_gpu (fully_fused_mlp.cu:757-762): needs the activated network output y
        assert y is not None, "mlp_backward needs the forward output when the network has an output activation"
        g = activation_backward(g, y, cfg.output_activation)
    for k in range(len(ws) - 1, -1, -1):
        inp = hidden[k - 1] if k > 0 else x.astype(np.float16)
        grads[k] = _dw_splitk16(g, inp) if isinstance(dw_acc, str) and dw_acc == "splitk16" else matmul_acc(np.ascontiguousarray(g.T), inp, dw_acc).astype(np.float64)
        z = matmul_acc(g, ws[k], acc)
        g = z.astype(np.float16)
        if k > 0:
            g = activation_backward(g, hidden[k - 1], cfg.activation)
    return np.concatenate([gr.reshape(-1) for gr in grads]), g


# --------------------------------------------------------------------------------------------------
# Losses (losses/l2.h:40-75, losses/relative_l2.h:40-75)
# --------------------------------------------------------------------------------------------------
def loss_l2(pred: np.ndarray, target: np.ndarray, n_out: int, loss_scale: float = LOSS_SCALE, relative: bool = False):
    """pred fp16 [B, out_pad], target fp32 [B, n_out] -> (mean loss float, dL/dpred fp16 [B, out_pad])."""
    B, stride = pred.shape
    n_total = F32(B * n_out)
    p = pred[:, :n_out].astype(np.float32)
    diff = (p - target.astype(np.float32)).astype(np.float32)
    denom = (p * p + F32(0.01)).astype(np.float32) if relative else np.ones_like(p)
    values = (diff * diff / denom / n_total).astype(np.float32)
    grad = (F32(2) * diff / denom).astype(np.float32)
    g = np.zeros((B, stride), dtype=np.float16)
    g[:, :n_out] = (F32(loss_scale) * grad / n_total).astype(np.float16)
    return float(values.astype(np.float64).sum()), g


# --------------------------------------------------------------------------------------------------
# Adam (optimizers/adam.h:47-119; hyper-parameter defaults :312-326)
# --------------------------------------------------------------------------------------------------
@dataclass
class AdamConfig:
    learning_rate: float = 1e-3
    beta1: float = 0.9
    beta2: float = 0.999
    epsilon: float = 1e-8
    l2_reg: float = 1e-8
    relative_decay: float = 0.0
    absolute_decay: float = 0.0
    adabound: bool = False


def adam_step(cfg: AdamConfig, n_matrix: int, master: np.ndarray, grads_h: np.ndarray, m: np.ndarray, v: np.ndarray,
              steps: np.ndarray, loss_scale: float = LOSS_SCALE):
    """In-place Adam on fp32 master weights; returns the fp16 copy.  Non-matrix (grid) params with gradient == 0 are
    skipped entirely and every parameter keeps its own step counter (adam.h:75-91,103-105)."""
    n = master.shape[0]
    grad = (grads_h.astype(np.float32) / F32(loss_scale)).astype(np.float32)
    is_mat = np.arange(n) < n_matrix
    active = is_mat | (grad != 0)
    grad = np.where(is_mat, (grad + F32(cfg.l2_reg) * master).astype(np.float32), grad)
    b1, b2 = F32(cfg.beta1), F32(cfg.beta2)
    m_new = (b1 * m + (F32(1) - b1) * grad).astype(np.float32)
    v_new = (b2 * v + (F32(1) - b2) * (grad * grad).astype(np.float32)).astype(np.float32)
    t = (steps + active.astype(np.uint32)).astype(np.uint32)
    tf = np.maximum(t, 1).astype(np.float32)
    lr = (F32(cfg.learning_rate) * np.sqrt(F32(1) - np.power(b2, tf, dtype=np.float32)) / (F32(1) - np.power(b1, tf, dtype=np.float32))).astype(np.float32)
    lo, hi = (F32(0.0), F32(np.finfo(np.float32).max))
    eff = np.minimum(np.maximum(lr / (np.sqrt(v_new) + F32(cfg.epsilon)), lo), hi).astype(np.float32)
    decayed = master  # weight_decay with zero relative/absolute decay is the identity (common_device.h:870-873)
    new_w = (decayed - eff * m_new).astype(np.float32)
    master[:] = np.where(active, new_w, master)
    m[:] = np.where(active, m_new, m)
    v[:] = np.where(active, v_new, v)
    steps[:] = t
    return master.astype(np.float16)


# --------------------------------------------------------------------------------------------------
# Parameter initialisation (trainer.h:50-87; network_with_input_encoding.h:124-130; fully_fused_mlp.cu:866-891;
# gpu_matrix.h:284-299 xavier on the host; grid.h:1059-1062 + random.h:39-71 uniform(-1e-4,1e-4) on the device)
# --------------------------------------------------------------------------------------------------
def random_uniform(rng: Pcg32, n: int) -> np.ndarray:
    """generate_random_uniform<float>(stream, rng, n, out) with lower = 0, upper = 1 (random.h:39-71): thread i of
    n_threads = 128 * ceil(ceil(n/4)/128) advances the generator by 4 i and writes element i + n_threads * j with its j-th
    draw; afterwards the host generator is advanced by n.  Used by samples/benchmarks to draw training batches."""
    n_threads = next_multiple((n + 3) // 4, 128)
    draws = _uint_to_float01(pcg32_stream(rng.copy(), 4 * n_threads)).reshape(n_threads, 4)
    flat = draws.T.reshape(-1)[:n].copy()  # val * (1 - 0) + 0 is exact
    rng.advance(n)
    return flat


def synth_target(x: np.ndarray, n_out: int) -> np.ndarray:
    """The analytic training target of oracle/ref_driver.cu (synth_target kernel): fp32 throughout."""
    n, n_in = x.shape
    s = np.zeros(n, np.float32)
    for d in range(n_in):
        s = (s + x[:, d] * F32(1.0 + 0.5 * d)).astype(np.float32)
    t = np.empty((n, n_out), np.float32)
    for c in range(n_out):
        t[:, c] = F32(0.5) + F32(0.5) * np.sin((F32(6.2831853) * F32(c + 1) * s).astype(np.float32), dtype=np.float32)
    return t


def init_params(mlp: MlpConfig, grid: GridConfig | None, seed: int = 1337, trainer_seeding: bool = True) -> np.ndarray:
    """fp32 [P_mlp + P_grid], MLP first.  trainer_seeding=True reproduces Trainer (seed_seq -> pcg32);
    False reproduces cpp::Module::initialize_params (pcg32{seed} directly, src/cpp_api.cu:124-127)."""
    rng = Pcg32(seed_seq_first(seed)) if trainer_seeding else Pcg32(seed)
    parts = []
    for (rows, cols) in mlp.shapes():
        scale = F32(np.sqrt(F32(6.0) / F32(cols + rows)))
        u = _uint_to_float01(pcg32_stream(rng, rows * cols))
        parts.append((u * F32(2.0) * scale - scale).astype(np.float32))
    if grid is not None:
        n = grid.n_params
        # generate_random_kernel: thread i advances by 4i and writes element i + n_threads*j with its j-th draw,
        # n_threads = 128 * ceil(ceil(n/4)/128)   (random.h:39-63, common_host.h n_blocks_linear / N_THREADS_LINEAR=128)
        n_threads = next_multiple((n + 3) // 4, 128)
        draws = _uint_to_float01(pcg32_stream(rng.copy(), 4 * n_threads)).reshape(n_threads, 4)
        flat = draws.T.reshape(-1)[:n]  # element idx = i + n_threads*j  <- draws[i, j]
        rng.advance(n)
        lo, hi = F32(-1e-4), F32(1e-4)
        # the device lambda `val * (upper - lower) + lower` is contracted to one fmaf by nvcc (default -fmad=true)
        rangef = F32(hi - lo)
        parts.append((flat.astype(np.float64) * np.float64(rangef) + np.float64(lo)).astype(np.float32))
    return np.concatenate(parts)


# --------------------------------------------------------------------------------------------------
# OneBlob encoding (encodings/oneblob.h:47-160; quartic kernel common_device.h:905-920) - BASELINE config #2
# --------------------------------------------------------------------------------------------------
def quartic_cdf(x: np.ndarray, inv_radius: float) -> np.ndarray:
    """common_device.h:915-920, fp32."""
    u = (x.astype(np.float32) * F32(inv_radius)).astype(np.float32)
    u2 = (u * u).astype(np.float32)
    u4 = (u2 * u2).astype(np.float32)
    poly = (F32(1.0) - F32(F32(2.0) / F32(3.0)) * u2 + F32(F32(1.0) / F32(5.0)) * u4).astype(np.float32)
    v = (F32(15.0 / 16.0) * u * poly + F32(0.5)).astype(np.float32)
    return np.maximum(F32(0.0), np.minimum(F32(1.0), v)).astype(np.float32)


def quartic_pdf(x: np.ndarray, inv_radius: float) -> np.ndarray:
    """quartic_cdf_deriv, common_device.h:905-913."""
    u = (x.astype(np.float32) * F32(inv_radius)).astype(np.float32)
    t = np.maximum(F32(1.0) - u * u, F32(0.0)).astype(np.float32)
    return (F32(15.0 / 16.0) * t * t * F32(inv_radius)).astype(np.float32)


def oneblob_shuffle_source(n_out: int, n_bins: int) -> np.ndarray:
    """Column whose LEFT-edge CDF a column takes as its RIGHT-edge CDF (one_blob_subwarp_aligned, oneblob.h:47-70):
    `__shfl_sync(0xffffffff, left_cdf, bin_index + 1, n_bins)` in a block of (n_out, ceil(128/n_out)) threads.
    For n_bins <= 32 that is the next bin of the same dimension, wrapping inside the dimension.  For n_bins = 64 the
    shuffle width exceeds the warp: the hardware then uses lane (bin + 1) mod 32 of the SAME warp (segment mask 0), so
    bins 31 and 63 read bins 0 and 32 - the reference's actual output, which parity follows."""
    j = np.arange(n_out)
    b = j % n_bins
    if n_bins <= 32:
        return (j - b) + (b + 1) % n_bins
    return (j & ~31) | ((b + 1) & 31)


def oneblob_forward(x: np.ndarray, n_bins: int, width: int | None = None) -> np.ndarray:
    """x fp32 [B, D] -> fp16 [B, width]; columns >= D * n_bins are 1 (oneblob.h:131-133).  AoS kernel (oneblob.h:84-96)."""
    B, D = x.shape
    n_out = D * n_bins
    width = width or next_multiple(n_out, 16)
    j = np.arange(n_out)
    b = (j % n_bins).astype(np.float32)
    xv = x.astype(np.float32)[:, j // n_bins]
    lb = (b / F32(n_bins)).astype(np.float32)[None, :]  # scalbnf(bin, -log2 n_bins): exact
    d = (lb - xv).astype(np.float32)
    left = (quartic_cdf(d, n_bins) + quartic_cdf((d - F32(1.0)).astype(np.float32), n_bins)).astype(np.float32)
    left = (left + quartic_cdf((d + F32(1.0)).astype(np.float32), n_bins)).astype(np.float32)
    right = left[:, oneblob_shuffle_source(n_out, n_bins)].copy()
    right[:, (j % n_bins) == n_bins - 1] += F32(1.0)
    out = np.ones((B, width), np.float16)
    out[:, :n_out] = (right - left).astype(np.float16)
    return out


def oneblob_backward_input(x: np.ndarray, n_bins: int, dy: np.ndarray) -> np.ndarray:
    """dL/dx fp32 [B, D] from dL/dy fp16 [B, >= D * n_bins] (kernel_one_blob_backward, oneblob.h:105-141)."""
    B, D = x.shape
    xv = x.astype(np.float32)

    def wrapped(boundary):
        d = (F32(boundary) - xv).astype(np.float32)
        return (quartic_pdf(d, n_bins) + quartic_pdf(d - F32(1.0), n_bins) + quartic_pdf(d + F32(1.0), n_bins)).astype(np.float32)

    res = np.zeros((B, D), np.float32)
    left = wrapped(0.0)
    for k in range(n_bins):
        right = wrapped((k + 1) / n_bins)
        res += dy[:, np.arange(D) * n_bins + k].astype(np.float32) * (left - right)
        left = right
    return res


# --------------------------------------------------------------------------------------------------
# Second order of the grid encoding (encodings/grid.h:351-650): the backward pass of dL/dx = sum_k dL/dy_k * dy_k/dx,
# i.e. the gradients of  S = sum_{i,d} dL_ddLdx[i,d] * dL/dx[i,d]  with respect to dL/dy, the table and x.
# Used by eikonal / SDF losses through tcnn.Encoding (bindings/torch/tinycudann/modules.py:120-160).
# --------------------------------------------------------------------------------------------------
def smoothstep_2nd_derivative(v: np.ndarray) -> np.ndarray:
    """common_device.h:809-811."""
    return (F32(6.0) - F32(12.0) * v.astype(np.float32)).astype(np.float32)


def _level_frac(cfg: GridConfig, x: np.ndarray, l: int):
    """(interpolation weight w, dw/dfrac, d2w/dfrac2, cell) of level l - pos_fract with three outputs, common_device.h:825-839."""
    frac, cell = pos_fract(x, F32(cfg.scales[l]))
    if cfg.smooth:
        return smoothstep(frac), smoothstep_derivative(frac), smoothstep_2nd_derivative(frac), cell
    return frac, np.ones_like(frac), np.zeros_like(frac), cell


def grid_dy_dx(cfg: GridConfig, x: np.ndarray, table: np.ndarray) -> np.ndarray:
    """dy_k/dx_d as kernel_grid stores it (grid.h:171-211): fp32 [B, L*F, D]."""
    B, D = x.shape
    F = cfg.n_features_per_level
    t = table.astype(np.float16).reshape(-1, F)
    out = np.zeros((B, cfg.n_features, D), np.float32)
    for l in range(cfg.active_levels):
        scale, res = F32(cfg.scales[l]), int(cfg.resolutions[l])
        w, dw, _, cell = _level_frac(cfg, x, l)
        size, base = cfg.level_size(l), int(cfg.offsets[l])
        for gd in range(D):
            for idx in range(1 << (D - 1)):
                weight = np.full(B, scale, np.float32)
                pg = np.empty((B, D), np.uint32)
                for nd in range(D - 1):
                    dim = nd + 1 if nd >= gd else nd
                    if (idx & (1 << nd)) == 0:
                        weight = (weight * (F32(1.0) - w[:, dim])).astype(np.float32); pg[:, dim] = cell[:, dim]
                    else:
                        weight = (weight * w[:, dim]).astype(np.float32); pg[:, dim] = cell[:, dim] + np.uint32(1)
                pg[:, gd] = cell[:, gd]
                vl = t[base + grid_index(pg, size, res, cfg.hashed, cfg.hash).astype(np.int64)].astype(np.float32)
                pg[:, gd] = cell[:, gd] + np.uint32(1)
                vr = t[base + grid_index(pg, size, res, cfg.hashed, cfg.hash).astype(np.int64)].astype(np.float32)
                out[:, l * F:(l + 1) * F, gd] += (weight[:, None] * (vr - vl)).astype(np.float32) * dw[:, gd, None]
    return out


def grid_bwd_bwd_dLddLdy(cfg: GridConfig, x: np.ndarray, table: np.ndarray, dL_ddLdx: np.ndarray) -> np.ndarray:
    """kernel_grid_backward_input_backward_dLdoutput (grid.h:624-650): fp16 [B, L*F], result_k = sum_d dy_k/dx_d * dL_ddLdx_d."""
    dydx = grid_dy_dx(cfg, x, table)
    return np.einsum("bkd,bd->bk", dydx.astype(np.float32), dL_ddLdx.astype(np.float32)).astype(np.float16)


def grid_bwd_bwd_grid(cfg: GridConfig, x: np.ndarray, dy: np.ndarray, dL_ddLdx: np.ndarray) -> np.ndarray:
    """kernel_grid_backward_input_backward_grid (grid.h:351-451): float64 [n_params]; per corner pair along grad_dim the table
    entries receive -/+ (half)weight * dL/dy with weight = scale * dL_ddLdx[gd] * dw[gd] * prod(other interpolation weights)."""
    B, D = x.shape
    F = cfg.n_features_per_level
    g = np.zeros((int(cfg.offsets[-1]), F), np.float64)
    for l in range(cfg.active_levels):
        scale, res = F32(cfg.scales[l]), int(cfg.resolutions[l])
        w, dw, _, cell = _level_frac(cfg, x, l)
        size, base = cfg.level_size(l), int(cfg.offsets[l])
        gl = dy[:, l * F:(l + 1) * F].astype(np.float16).astype(np.float32)
        for gd in range(D):
            grad_in = (scale * dL_ddLdx[:, gd].astype(np.float32) * dw[:, gd]).astype(np.float32)
            for idx in range(1 << (D - 1)):
                weight = grad_in.copy()
                pg = np.empty((B, D), np.uint32)
                for nd in range(D - 1):
                    dim = nd + 1 if nd >= gd else nd
                    if (idx & (1 << nd)) == 0:
                        weight = (weight * (F32(1.0) - w[:, dim])).astype(np.float32); pg[:, dim] = cell[:, dim]
                    else:
                        weight = (weight * w[:, dim]).astype(np.float32); pg[:, dim] = cell[:, dim] + np.uint32(1)
                for side, sign in ((0, -1.0), (1, 1.0)):
                    pg[:, gd] = cell[:, gd] + np.uint32(side)
                    wh = (F32(sign) * weight).astype(np.float16).astype(np.float32)
                    prod = (wh[:, None] * gl).astype(np.float16)  # (GRAD_T)weight * grad: half multiply
                    np.add.at(g, base + grid_index(pg, size, res, cfg.hashed, cfg.hash).astype(np.int64), prod.astype(np.float64))
    return g.reshape(-1)


def grid_bwd_bwd_input(cfg: GridConfig, x: np.ndarray, table: np.ndarray, dy: np.ndarray, dL_ddLdx: np.ndarray) -> np.ndarray:
    """kernel_grid_backward_input_backward_input (grid.h:453-622): fp32 [B, D], the Hessian of the interpolation contracted with
    dL/dy and dL_ddLdx - diagonal part (Smoothstep only: d2w/dfrac2) and the cross terms between grad_dim and the other dims."""
    B, D = x.shape
    F = cfg.n_features_per_level
    t = table.astype(np.float16).reshape(-1, F).astype(np.float32)
    out = np.zeros((B, D), np.float32)
    v = dL_ddLdx.astype(np.float32)
    for l in range(cfg.active_levels):
        scale, res = F32(cfg.scales[l]), int(cfg.resolutions[l])
        w, dw, d2w, cell = _level_frac(cfg, x, l)
        size, base = cfg.level_size(l), int(cfg.offsets[l])
        gl = dy[:, l * F:(l + 1) * F].astype(np.float16).astype(np.float32)

        def calc(pg, weight):
            vals = t[base + grid_index(pg, size, res, cfg.hashed, cfg.hash).astype(np.int64)]
            return ((vals * gl).sum(axis=1) * weight).astype(np.float32)

        diag = (scale * scale * v * d2w).astype(np.float32)
        other = (scale * scale * v * dw).astype(np.float32)
        for gd in range(D):
            acc = np.zeros(B, np.float32)
            for idx in range(1 << (D - 1)):
                if cfg.smooth:
                    weight = diag[:, gd].copy()
                    pg = np.empty((B, D), np.uint32)
                    for nd in range(D - 1):
                        dim = nd + 1 if nd >= gd else nd
                        if (idx & (1 << nd)) == 0:
                            weight = weight * (F32(1.0) - w[:, dim]); pg[:, dim] = cell[:, dim]
                        else:
                            weight = weight * w[:, dim]; pg[:, dim] = cell[:, dim] + np.uint32(1)
                    pg[:, gd] = cell[:, gd]; acc += calc(pg, -weight)
                    pg[:, gd] = cell[:, gd] + np.uint32(1); acc += calc(pg, weight)
                for og in range(D - 1):
                    rog = og + 1 if og >= gd else og  # the other gradient dim
                    weight = (other[:, rog] * dw[:, gd]).astype(np.float32)
                    pg = np.empty((B, D), np.uint32)
                    for nd in range(D - 1):
                        dim = nd + 1 if nd >= rog else nd
                        if (idx & (1 << nd)) == 0:
                            weight = weight * (F32(1.0) - w[:, dim]) if dim != gd else -weight
                            pg[:, dim] = cell[:, dim]
                        else:
                            if dim != gd:
                                weight = weight * w[:, dim]
                            pg[:, dim] = cell[:, dim] + np.uint32(1)
                    pg[:, rog] = cell[:, rog]; acc += calc(pg, -weight)
                    pg[:, rog] = cell[:, rog] + np.uint32(1); acc += calc(pg, weight)
            out[:, gd] += acc
    return out
