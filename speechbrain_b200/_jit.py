"""``torch.jit.trace`` support for modules whose arithmetic happens behind ctypes.

The reference's unit tests trace every feature module
(tests/unittests/test_features.py:14,39,56,66,97,107).  A ctypes call is invisible to
the tracer - the recorded graph would return an uninitialised ``torch.empty`` - so under
tracing each module ``forward`` is routed through ONE registered torch operator,
``sbfe::module_call(Tensor[] inputs, int handle) -> Tensor``: the tracer records that node,
and replaying the graph re-enters the same Python ``forward`` (and therefore the same
libsbfe.so kernels) with the traced constants (non-tensor arguments, module configuration)
frozen, which is what tracing means.  Outside tracing nothing changes: ``forward`` calls the
C ABI directly, no dispatcher hop.

The operator is defined in Python, so a traced graph is valid in the process that traced
it (like any Python-defined custom op); it is not a serialisation format.
"""
import functools
import threading
from typing import List

import torch

__all__ = ["traceable"]

_registry = {}
_lock = threading.Lock()
_local = threading.local()


@torch.library.custom_op("sbfe::module_call", mutates_args=())
def module_call(inputs: List[torch.Tensor], handle: int) -> torch.Tensor:
    fn = _registry[handle]
    _local.inside = getattr(_local, "inside", 0) + 1
    try:
        out = fn(*inputs)
    finally:
        _local.inside -= 1
    # an operator result may not alias an input (in-place modes return their argument)
    if any(out is t or out.data_ptr() == t.data_ptr() for t in inputs if t.numel()):
        out = out.clone()
    return out


def _register(module, key, fn):
    """One handle per (module, frozen constants): tracing the same call twice must give the same
    graph (torch.jit.trace's own sanity check re-traces and compares)."""
    cache = module.__dict__.setdefault("_sbfe_trace_handles", {})
    handle = cache.get(key)
    with _lock:
        if handle is None:
            handle = cache[key] = len(_registry)
        _registry[handle] = fn
    return handle


def _tracing():
    return torch._C._get_tracing_state() is not None and not getattr(_local, "inside", 0)


def plain_eager():
    """True outside ``torch.jit.trace`` and outside a replay of a traced graph: only then can
    python attributes on a returned tensor reach the next module (the tile-record hand-over)."""
    return torch._C._get_tracing_state() is None and not getattr(_local, "inside", 0)


def traceable(inplace_arg=None):
    """Decorate ``forward(self, *tensors, **constants)``.  ``inplace_arg`` = index of the tensor
    argument the eager ``forward`` overwrites and returns (None: returns a fresh tensor, or
    decided per call by a callable ``inplace_arg(self) -> index | None``)."""

    def deco(forward):
        @functools.wraps(forward)
        def wrapper(self, *args, **kwargs):
            if not _tracing():
                return forward(self, *args, **kwargs)
            tensors = [a for a in args if isinstance(a, torch.Tensor)]
            if len(tensors) != len(args):
                # non-tensor positionals become trace constants
                slots = [i for i, a in enumerate(args) if isinstance(a, torch.Tensor)]
                frozen = list(args)

                def call(*ts):
                    full = list(frozen)
                    for i, t in zip(slots, ts):
                        full[i] = t
                    return forward(self, *full, **kwargs)
            else:
                def call(*ts):
                    return forward(self, *ts, **kwargs)
            idx = inplace_arg(self) if callable(inplace_arg) else inplace_arg
            key = (forward.__qualname__, idx,
                   repr([None if isinstance(a, torch.Tensor) else a for a in args]),
                   repr(sorted((k, None if isinstance(v, torch.Tensor) else v)
                               for k, v in kwargs.items())))
            if idx is not None:
                # keep the aliasing contract visible to the tracer: compute, then copy_ into x
                target = args[idx]
                shadow = list(tensors)
                pos = [i for i, a in enumerate(args) if isinstance(a, torch.Tensor)].index(idx)

                def call_clone(*ts, _call=call, _pos=pos):
                    ts = list(ts)
                    ts[_pos] = ts[_pos].clone()
                    return _call(*ts)
                out = torch.ops.sbfe.module_call(shadow, _register(self, key, call_clone))
                return target.copy_(out)
            return torch.ops.sbfe.module_call(tensors, _register(self, key, call))

        return wrapper

    return deco
