"""A few values remembered per tensor OBJECT (identity + in-place version counter).

The reference reads device-resident metadata on the host in every layer of every frame (the Len_in assertion,
ms_deform_attn.py:62; H and W of every level, deformable_transformer.py:311-314): one device->host sync each,
and illegal while a CUDA graph is being captured.  The same few ``spatial_shapes`` tensors come back call after
call (one per pyramid; two alternating ones in the multi-frame configuration: the encoder's levels and the
decoder's twice as many), so the derived values are kept for the last few tensor objects seen.  A tensor that
was modified in place (version counter) or garbage-collected and replaced (weak reference) misses.
Tensors created under ``torch.inference_mode()`` keep no version counter (reading it raises); they cannot be
modified in place outside inference mode, so for them identity + weak reference is the whole key.
"""
import weakref
from collections import OrderedDict


def _version(tensor):
    return None if tensor.is_inference() else tensor._version


class PerTensorCache:
    def __init__(self, maxsize=8):
        self.maxsize = maxsize
        self._entries = OrderedDict()       # id(tensor) -> (weakref, version, extra key, value)

    def get(self, tensor, extra=None):
        e = self._entries.get(id(tensor))
        if e is not None and e[0]() is tensor and e[1] == _version(tensor) and e[2] == extra:
            return e[3]
        return None

    def put(self, tensor, value, extra=None):
        self._entries[id(tensor)] = (weakref.ref(tensor), _version(tensor), extra, value)
        self._entries.move_to_end(id(tensor))
        while len(self._entries) > self.maxsize:
            self._entries.popitem(last=False)
        return value

    def __len__(self):
        return len(self._entries)
