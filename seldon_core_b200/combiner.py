"""Combiner drop-in: elementwise mean over the child outputs, on the GPU.

Mirrors ``examples/combiners/mnist_combiner/MnistCombiner.py:8-11`` (``aggregate(Xs, features_names)``) generalised
to N children with the checks of ``engine/.../AverageCombinerUnit.java:35-79`` (no inputs, non 2-D data and shape
mismatches are errors; the messages are the engine's).
"""
import ctypes

import numpy as np

from . import _cabi
from .errors import SeldonB200Error


class MeanCombinerB200(object):
    def __init__(self, device=None, **kwargs):
        self._device = device
        self._wire = None

    def __getstate__(self):
        return {"_device": None}

    def __setstate__(self, state):
        self._device = None
        self._wire = None

    def aggregate_raw(self, msgs):
        """The hook the wrapper tries first (seldon_methods.py:268-272, user_model.py:77-78).  `msgs`: a list of
        serialised SeldonMessages (bytes / `wire.HostBuf` / uint8 CPU tensors), the bytes of a serialised
        SeldonMessageList, or a protobuf SeldonMessageList object.  Every child's payload is decoded on the device,
        the mean is taken there, and the response (payload kind and puid of the first message, names t:0.., empty
        meta - construct_response(user_model, False, request.seldonMessages[0], ...), seldon_methods.py:296) is
        encoded on the device.  Returns a `wire.HostBuf`, or a message object when given one."""
        import torch

        from . import wire

        if not torch.cuda.is_available():
            raise SeldonB200Error("no CUDA device: seldon_core_b200 has no CPU fallback", status_code=500)
        proto_cls = None
        if hasattr(msgs, "seldonMessages"):
            items = list(msgs.seldonMessages)
            proto_cls = type(items[0]) if items else None
        elif isinstance(msgs, (list, tuple)):
            items = list(msgs)
        else:
            items = None
        dev = torch.device(self._device or ("cuda:%d" % torch.cuda.current_device()))
        if self._wire is None:
            self._wire = wire.WireStage(dev)
        st = self._wire
        mats, first = [], None
        if items is None:  # one buffer holding a SeldonMessageList
            host = st.host_bytes(msgs)
            for off, ln in wire.split_message_list(host.numpy()):
                info = wire.parse_seldon_message(host.numpy(), off, off + ln)
                first = first or info
                mats.append(st.decode_matrix(host, info))
        else:
            for m in items:
                host = st.host_bytes(m)
                info = wire.parse_seldon_message(host.numpy())
                first = first or info
                mats.append(st.decode_matrix(host, info))
        if not mats:
            raise SeldonB200Error("Combiner received no inputs")
        for x in mats:
            if x.shape[0] != mats[0].shape[0]:
                raise SeldonB200Error("Expected batch length %d but found %d" % (mats[0].shape[0], x.shape[0]))
            if x.shape[1] != mats[0].shape[1]:
                raise SeldonB200Error("Expected batch length %d but found %d" % (mats[0].shape[1], x.shape[1]))
        out = self.aggregate_device(mats)
        names = ["t:{}".format(i) for i in range(out.shape[1])]
        resp = wire.build_message(st, first["kind"], names, out, puid=first["puid"])
        if proto_cls is not None:
            return proto_cls.FromString(resp.tobytes())
        return resp

    def aggregate(self, Xs, features_names=None):
        """numpy in -> numpy float64 out (what the wrapper expects); CUDA tensors in -> CUDA float32 tensor out."""
        import torch

        if len(Xs) == 0:
            raise SeldonB200Error("Combiner received no inputs")
        on_device = all(isinstance(x, torch.Tensor) and x.is_cuda for x in Xs)
        shape0 = tuple(Xs[0].shape)
        if len(shape0) != 2:
            raise SeldonB200Error("Combiner received data that is not 2 dimensional")
        for x in Xs:
            if len(x.shape) != 2:
                raise SeldonB200Error("Combiner received data that is not 2 dimensional")
            if x.shape[0] != shape0[0]:
                raise SeldonB200Error("Expected batch length %d but found %d" % (shape0[0], x.shape[0]))
            if x.shape[1] != shape0[1]:
                raise SeldonB200Error("Expected batch length %d but found %d" % (shape0[1], x.shape[1]))
        if not torch.cuda.is_available():
            raise SeldonB200Error("no CUDA device: seldon_core_b200 has no CPU fallback", status_code=500)
        dev = Xs[0].device if on_device else torch.device(self._device or ("cuda:%d" % torch.cuda.current_device()))
        if on_device:
            ins = [x.to(torch.float32).contiguous() for x in Xs]
        else:
            ins = [torch.from_numpy(np.ascontiguousarray(np.asarray(x), dtype=np.float32)).to(dev, non_blocking=True)
                   for x in Xs]
        out = self.aggregate_device(ins)
        return out if on_device else out.double().cpu().numpy()

    @staticmethod
    def aggregate_device(ins):
        """ins: list of contiguous CUDA float32 tensors of one shape.  Returns their mean (CUDA float32)."""
        import torch

        out = torch.empty_like(ins[0])
        ptrs = (ctypes.c_void_p * len(ins))(*[t.data_ptr() for t in ins])
        with torch.cuda.device(out.device):
            _cabi.check(_cabi.lib().sb_combiner_mean(ptrs, len(ins), out.numel(), out.data_ptr(),
                                                     torch.cuda.current_stream().cuda_stream))
        return out
