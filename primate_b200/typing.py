"""Keyword-argument plumbing (mirror of src/primate/typing.py:5-14)."""
import inspect
from typing import Callable


def _param_names(fun: Callable) -> set:
	return set(inspect.signature(fun).parameters.keys())


def restrict_kwargs(fun: Callable, kwargs) -> dict:
	"""Keeps only the entries of `kwargs` that `fun` accepts."""
	names = _param_names(fun)
	return {key: val for key, val in kwargs.items() if key in names}


def setdiff_kwargs(f: Callable, kwargs) -> dict:
	"""Keeps only the entries of `kwargs` that `f` does NOT accept."""
	names = _param_names(f)
	return {key: val for key, val in kwargs.items() if key not in names}
