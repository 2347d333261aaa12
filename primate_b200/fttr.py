"""Forward three-term-recurrence quadrature weights on the device (API mirror of src/primate/fttr.py:5-29)."""
import numpy as np

from . import _ffi, engine


def fttr(theta: np.ndarray, alpha: np.ndarray, beta: np.ndarray, k: int, weights: np.ndarray) -> None:
	"""weights[i] = 1 / (mu_0 * sum_j p_j(theta_i)^2) for the orthonormal polynomials p_j of the Jacobi matrix
	(alpha, beta) with beta[0] = 0 and mu_0 = sum |theta[:k]|; written in place like the reference kernel."""
	torch = engine._torch()
	n = len(alpha)
	assert k <= len(theta) and len(beta) >= n
	th = torch.from_numpy(np.ascontiguousarray(theta[:k], dtype=np.float64)[None, :].copy()).cuda()
	a = torch.from_numpy(np.ascontiguousarray(alpha, dtype=np.float64)[None, :].copy()).cuda()
	b = torch.from_numpy(np.ascontiguousarray(beta[:n], dtype=np.float64)[None, :].copy()).cuda()
	tau = torch.empty((1, k), dtype=torch.float64, device="cuda")
	_ffi.check(_ffi.lib().slq_fttr(th.data_ptr(), a.data_ptr(), b.data_ptr(), _ffi.SLQ_F64, 1, int(k), int(n), int(n), tau.data_ptr(),
	                               engine._stream_ptr()))  # fmt: skip
	weights[:k] = tau.cpu().numpy()[0]
