"""The three rlpyt helpers the hot path touches.

When rlpyt is importable (the reference checkout is on ``sys.path``) its own classes are used, so
``RSSMState`` is *the* rlpyt ``namedarraytuple`` type and interoperates with rlpyt buffers.  On a box
without the reference (the GPU box) the minimal re-implementations below provide the same
behaviour for what the path needs: field-wise indexing, ``buffer_method`` / ``buffer_func`` over
nested tuples, and leading-dimension bookkeeping.
Behaviour follows rlpyt/rlpyt/utils/collections.py:16-135, buffer.py:156-205, tensor.py:49-86.
"""
from __future__ import annotations

import collections

try:  # pragma: no cover - depends on the environment
    from rlpyt.utils.collections import namedarraytuple  # type: ignore
    from rlpyt.utils.buffer import buffer_method, buffer_func  # type: ignore
    from rlpyt.utils.tensor import infer_leading_dims, restore_leading_dims  # type: ignore

    HAVE_RLPYT = True
except Exception:  # rlpyt absent: local equivalents
    HAVE_RLPYT = False

    def namedarraytuple(typename, field_names):
        base = collections.namedtuple(typename, field_names)

        def getitem(self, loc):
            # an index applies to every field (None fields pass through)
            return type(self)(*(None if f is None else f[loc] for f in self))

        def setitem(self, loc, value):
            if not (isinstance(value, tuple) and getattr(value, "_fields", None) == self._fields):
                value = tuple(None if f is None else value for f in self)
            for f, v in zip(self, value):
                if f is not None:
                    f[loc] = v

        def items(self):
            for k, v in zip(self._fields, self):
                yield k, v

        cls = type(typename, (base,), {
            "__slots__": (), "__getitem__": getitem, "__setitem__": setitem, "items": items,
            "get": lambda self, i: tuple.__getitem__(self, i),
            "__contains__": lambda self, key: key in self._fields,
        })
        cls._is_namedarraytuple = True
        return cls

    def buffer_method(buffer_, method_name, *args, **kwargs):
        if isinstance(buffer_, tuple):
            contents = tuple(buffer_method(b, method_name, *args, **kwargs) for b in buffer_)
            return type(buffer_)(*contents) if hasattr(buffer_, "_fields") else contents
        if buffer_ is None:
            return None
        return getattr(buffer_, method_name)(*args, **kwargs)

    def buffer_func(buffer_, func, *args, **kwargs):
        if isinstance(buffer_, tuple):
            contents = tuple(buffer_func(b, func, *args, **kwargs) for b in buffer_)
            return type(buffer_)(*contents) if hasattr(buffer_, "_fields") else contents
        if buffer_ is None:
            return None
        return func(buffer_, *args, **kwargs)

    def infer_leading_dims(tensor, dim):
        lead_dim = tensor.dim() - dim
        assert lead_dim in (0, 1, 2)
        if lead_dim == 2:
            T, B = tensor.shape[:2]
        else:
            T = 1
            B = 1 if lead_dim == 0 else tensor.shape[0]
        return lead_dim, T, B, tensor.shape[lead_dim:]

    def restore_leading_dims(tensors, lead_dim, T=1, B=1):
        is_seq = isinstance(tensors, (tuple, list))
        ts = tensors if is_seq else (tensors,)
        if lead_dim == 2:
            ts = tuple(t.view((T, B) + t.shape[1:]) for t in ts)
        elif lead_dim == 0:
            assert B == 1 and T == 1
            ts = tuple(t.squeeze(0) for t in ts)
        return ts if is_seq else ts[0]
