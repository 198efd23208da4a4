"""Zipf key generator of the reference's succinct micro-benchmarks (BASELINE.json configs 1 and 3).

Restates `Zipf<uint64_t, double>` (benchmark/micro/succinct/zipf.cpp:12-100: rejection-inversion sampling,
Hoermann & Derflinger 1996) driven by `std::mt19937` through `std::uniform_real_distribution<double>`
(libstdc++: two 32-bit draws per double, low word first).  The reference seeds from `random_device`
(zipf_distribution.cpp:29-30); the workloads here fix the seed so that runs are repeatable.  This module only
makes benchmark INPUT (look-up keys); it is not part of the codec."""
from __future__ import annotations

import math

import numpy as np

_EPS = 1e-8


class Zipf:
    def __init__(self, n: int, q: float = 1.0, seed: int = 42):
        self.n, self.q = int(n), float(q)
        self.rs = np.random.RandomState(seed)  # MT19937 with init_genrand(seed) == std::mt19937(seed)
        self.h_x1 = self._H(1.5) - 1.0
        self.h_n = self._H(self.n + 0.5)
        self._buf = np.zeros(0, dtype=np.uint64)
        self._pos = 0

    # ---- zipf.cpp:50-94 ----
    @staticmethod
    def _expxm1bx(x: float) -> float:
        return math.expm1(x) / x if abs(x) > _EPS else 1.0 + x / 2.0 * (1.0 + x / 3.0 * (1.0 + x / 4.0))

    def _H(self, x: float) -> float:
        log_x = math.log(x)
        return self._expxm1bx((1.0 - self.q) * log_x) * log_x

    @staticmethod
    def _log1pxbx(x: float) -> float:
        return math.log1p(x) / x if abs(x) > _EPS else 1.0 - x * (0.5 - x * (1.0 / 3.0 - x * 0.25))

    def _H_inv(self, x: float) -> float:
        t = max(-1.0, x * (1.0 - self.q))
        return math.exp(self._log1pxbx(t) * x)

    def _h(self, x: float) -> float:
        return math.exp(-self.q * math.log(x))

    def _raw32(self) -> int:
        if self._pos >= len(self._buf):
            self._buf = self.rs.randint(0, 1 << 32, size=4096, dtype=np.uint64)
            self._pos = 0
        v = int(self._buf[self._pos])
        self._pos += 1
        return v

    def _uniform(self) -> float:
        # std::generate_canonical<double, 53>(mt19937): (lo + hi * 2^32) / 2^64, then scaled into [a, b)
        lo, hi = self._raw32(), self._raw32()
        r = (lo + hi * 4294967296.0) / 18446744073709551616.0
        if r >= 1.0:
            r = math.nextafter(1.0, 0.0)
        return r * (self.h_n - self.h_x1) + self.h_x1

    def __call__(self) -> int:  # zipf.cpp:32-42
        while True:
            u = self._uniform()
            x = self._H_inv(u)
            k = min(max(int(math.floor(x + 0.5)), 1), self.n)  # std::round: halves away from zero (x > 0)
            if u >= self._H(k + 0.5) - self._h(k):
                return k


def zipf_keys_scalar(n: int, q: float, count: int, seed: int = 42) -> np.ndarray:
    z = Zipf(n, q, seed)
    return np.fromiter((z() for _ in range(count)), dtype=np.int64, count=count)


def zipf_keys(n: int, q: float, count: int, seed: int = 42) -> np.ndarray:
    """The same key stream as `zipf_keys_scalar`, vectorised: every attempt of the rejection loop consumes exactly one
    uniform (two 32-bit draws), so evaluating the acceptance test on a whole block of uniforms and keeping the accepted
    ones in order reproduces the sequential sampler draw for draw."""
    z = Zipf(n, q, seed)
    one_minus_q = 1.0 - z.q

    def expxm1bx(x):
        small = np.abs(x) <= _EPS
        xs = np.where(small, 1.0, x)
        return np.where(small, 1.0 + x / 2.0 * (1.0 + x / 3.0 * (1.0 + x / 4.0)), np.expm1(xs) / xs)

    def H(x):
        log_x = np.log(x)
        return expxm1bx(one_minus_q * log_x) * log_x

    def log1pxbx(x):
        small = np.abs(x) <= _EPS
        xs = np.where(small, 1.0, x)
        return np.where(small, 1.0 - x * (0.5 - x * (1.0 / 3.0 - x * 0.25)), np.log1p(xs) / xs)

    out = np.zeros(0, dtype=np.int64)
    while len(out) < count:
        m = max(1024, int((count - len(out)) * 1.2))
        raw = z.rs.randint(0, 1 << 32, size=2 * m, dtype=np.uint64).astype(np.float64)
        r = (raw[0::2] + raw[1::2] * 4294967296.0) / 18446744073709551616.0
        r = np.where(r >= 1.0, np.nextafter(1.0, 0.0), r)
        u = r * (z.h_n - z.h_x1) + z.h_x1
        t = np.maximum(-1.0, u * one_minus_q)
        x = np.exp(log1pxbx(t) * u)
        k = np.clip(np.floor(x + 0.5), 1, z.n)
        ok = u >= H(k + 0.5) - np.exp(-z.q * np.log(k))
        acc = k[ok].astype(np.int64)
        need = count - len(out)
        if len(acc) >= need:
            # the sequential sampler would stop right after its `need`-th acceptance; later draws are simply unused
            acc = acc[:need]
        out = np.concatenate([out, acc])
    return out
