"""CPU restatement of NumPy's ziggurat normal sampler -- TEST INFRASTRUCTURE ONLY (see oracle/slq_oracle.py).

The reference's normal / sphere probes are `rng.standard_normal(out=W)` on a PCG64 Generator (src/primate/random.py:32-41);
that arithmetic lives in NumPy (pinned: numpy 2.3.5), numpy/random/src/distributions/distributions.c
`random_standard_normal` with the tables of ziggurat_constants.h.  Restated here on the raw PCG64 stream:

    r = next_uint64; idx = r & 0xff; r >>= 8; sign = r & 1; rabs = (r >> 1) & (2^52 - 1); x = rabs * wi[idx] (signed)
    rabs < ki[idx]                      -> x                                  (one raw draw, 99.3%)
    idx == 0 (tail)                     -> loop: xx = -log1p(-U1) / r, yy = -log1p(-U2); accept if 2 yy > xx^2
    else (wedge)                        -> accept x if (fi[idx-1] - fi[idx]) U + fi[idx] < exp(-x^2 / 2), else start over
    with U = (next_uint64 >> 11) * 2^-53.

Besides the values this returns the number of raw draws consumed (data dependent), which is what makes the stream
position of probe i a sequential quantity."""
from __future__ import annotations

import math
import os

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ZIG_R = 3.6541528853610087963519472518
ZIG_INV_R = 0.27366123732975827203338247596


def tables():
	z = np.load(os.path.join(os.path.dirname(HERE), "tests", "golden", "ziggurat_tables.npz"))
	return z["ki"], z["wi"], z["fi"]


def np_standard_normal(seed, n: int, tables=None, raw=None):
	"""(values[n], raw draws consumed) of `default_rng(seed).standard_normal(n)`, from the raw stream."""
	ki, wi, fi = tables if tables is not None else globals()["tables"]()
	if raw is None:
		bg = seed.bit_generator if isinstance(seed, np.random.Generator) else np.random.PCG64(seed)
		raw = np.random.PCG64()
		raw.state = bg.state  # a copy: the caller's generator is left alone
		raw = raw.random_raw(int(n * 1.03) + 4096)
	out = np.empty(n)
	t, i = 0, 0
	# vectorised fast path between slow samples
	idx_all = (raw & np.uint64(0xFF)).astype(np.int64)
	rr = raw >> np.uint64(8)
	sign_all = (rr & np.uint64(1)).astype(bool)
	rabs_all = (rr >> np.uint64(1)) & np.uint64(0x000FFFFFFFFFFFFF)
	x_all = rabs_all.astype(np.float64) * wi[idx_all]
	x_all = np.where(sign_all, -x_all, x_all)
	fast = rabs_all < ki[idx_all]
	slow_pos = np.nonzero(~fast)[0]
	sp = 0
	to_double = lambda u: float(u >> np.uint64(11)) * (1.0 / 9007199254740992.0)  # noqa: E731
	while i < n:
		while sp < len(slow_pos) and slow_pos[sp] < t:
			sp += 1
		nxt = int(slow_pos[sp]) if sp < len(slow_pos) else len(raw)
		m = min(nxt - t, n - i)
		if m > 0:
			out[i : i + m] = x_all[t : t + m]
			i += m
			t += m
			continue
		# a sample starting at a slow position
		while True:
			idx, rabs, x = int(idx_all[t]), int(rabs_all[t]), float(x_all[t])
			t += 1
			if rabs < int(ki[idx]):
				out[i] = x
				break
			if idx == 0:
				while True:
					xx = -ZIG_INV_R * math.log1p(-to_double(raw[t]))  # libm's log1p, as distributions.c calls (np.log1p may take a SIMD path)
					yy = -math.log1p(-to_double(raw[t + 1]))
					t += 2
					if yy + yy > xx * xx:
						break
				out[i] = -(ZIG_R + xx) if ((rabs >> 8) & 1) else ZIG_R + xx
				break
			u = to_double(raw[t])
			t += 1
			if (fi[idx - 1] - fi[idx]) * u + fi[idx] < math.exp(-0.5 * x * x):
				out[i] = x
				break
		i += 1
	return out, t
