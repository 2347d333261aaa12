"""primate_b200 -- a B200 (sm_100a) backend for the stochastic-Lanczos-quadrature path of peekxc/primate.

The sub-modules mirror primate's own (`lanczos`, `operators`, `trace`, `diagonal`, `integrate`, `tridiag`, `random`,
`estimators`, `stats`, `special`); they are imported on demand, as in the reference (src/primate/__init__.py:5-14 keeps
its imports commented out), so `import primate_b200` stays cheap and works on a box without a GPU.
"""
__version__ = "0.5.6+b200.2"  # tracks scikit-primate 0.5.6, the reference this backend is a drop-in for (src/primate/__init__.py:3)


def get_include() -> str:
	"""Directory of `slq.h`, the C header of libslq: what `primate.get_include()` is for the reference's C++ headers
	(src/primate/__init__.py:17-41).  Native code that wants to drive the engine without Python compiles against it
	and links `primate_b200/libslq.so`."""
	from ._build import get_include as _inc

	return _inc()


def get_library() -> str:
	"""Path of the built C-ABI shared library (built on first use; raises when it cannot be)."""
	from . import _ffi

	_ffi.lib()
	return _ffi.lib_path()
