"""Drop-in for the reference's online Mahalanobis outlier detector, computed on a B200.

Mirrors ``components/outlier-detection/mahalanobis/CoreMahalanobis.py`` of the reference: same constructor
parameters (L26), same methods (``predict`` L42, ``transform_input`` L54, ``send_feedback`` L154, ``tags`` L168,
``metrics`` L180) and the same public attributes (``threshold, n_components, max_n, n_stdev, start_clip, clip,
mean, C, n, nb_outliers, score, prediction``).  The arithmetic of ``_get_preds`` (L68-151) runs in libsb200.so
through the C ABI of ``include/sb200.h``; PyTorch only owns the device buffers and the stream.

Deliberate differences: scores are computed in fp32 (tolerance 1e-4 relative), the warm-up gate has
a rank guard, ``max_n=1`` is rejected (the reference raises ZeroDivisionError at L134), and at most 32 principal
components are supported.
"""
import ctypes
import threading

import numpy as np

from . import _cabi
from .errors import SeldonB200Error


class OutlierMahalanobisB200(object):
    def __init__(self, threshold=25, n_components=3, n_stdev=3, start_clip=50, max_n=-1, device=None):
        self.threshold = threshold
        self.n_components = n_components
        self.max_n = max_n
        self.n_stdev = n_stdev
        self.start_clip = start_clip
        self.nb_outliers = 0
        self._device = device
        self._h = None
        self._d = None
        self._ws = None
        self._lock = threading.RLock()
        self._restore = None
        self._score_t = None
        self._flag_t = None
        self._options = {}
        self._wire = None

    # ------------------------------------------------------------------ reference API
    def predict(self, X, feature_names=None):
        """Return outlier predictions (CoreMahalanobis.py:42-51): int array of 0/1, one per row."""
        return self._get_preds(X)

    def transform_input(self, X, feature_names=None):
        """Detect outliers, keep them for tags(), return the input unchanged (L54-65)."""
        self.prediction_meta = self._get_preds(X)
        return X

    def send_feedback(self, X, feature_names, reward, truth, routing=None):
        # L154-165.  `routing` is how the wrapper passes the request's routing (user_model.py:308-313); the detector has
        # no use for it.
        return []

    def tags(self):
        try:
            return {"outlier-predictions": self.prediction_meta.tolist()}  # L168-177
        except AttributeError:
            return None

    def metrics(self):
        """Six gauges averaged over the last batch; mutates nb_outliers like the reference (L180-192)."""
        with self._lock:
            nrows, n_flagged, score_sum = self._last_sums()
            self.nb_outliers += n_flagged
            n = self.n
            g = lambda key, value: {"type": "GAUGE", "key": key, "value": value}
            return [
                g("is_outlier", n_flagged / nrows),
                g("outlier_score", score_sum / nrows),
                g("nb_outliers", int(self.nb_outliers)),
                g("fraction_outliers", int(self.nb_outliers) / n),
                g("observation", n),
                g("threshold", self.threshold),
            ]

    # ------------------------------------------------------------------ raw hooks (bytes in -> bytes out)
    def predict_raw(self, msg, shard=None):
        """The hook the wrapper tries before anything else (seldon_methods.py:43-47, user_model.py:36-37): `msg` is a
        serialised SeldonMessage (bytes, a pinned `wire.HostBuf`, a uint8 CPU tensor), a protobuf SeldonMessage object
        (gRPC: re-serialised once) or the parsed REST dict.  The payload is decoded on the device, scored, and the
        response (flags as the request's payload kind, `names: []`, meta with tags and the six gauges, puid echoed -
        construct_response, utils.py:410-469) is encoded on the device.  Bytes-like in -> `wire.HostBuf` out (a view
        of this instance's pinned response buffer, valid until the next raw call); protobuf object in -> an object of
        the same class; dict in -> dict.  `shard=(ShardedMahalanobis, total_rows)` scores this rank's row block."""
        return self._raw(msg, False, shard)

    def transform_input_raw(self, msg, shard=None):
        """As predict_raw for the transformer mode (seldon_methods.py:124-128): the response echoes the request's
        `data` unchanged and carries the flags in `meta.tags["outlier-predictions"]` (CoreMahalanobis.py:54-65,
        168-177), packed on the device."""
        return self._raw(msg, True, shard)

    def _raw(self, msg, transform, shard):
        from . import wire

        if isinstance(msg, dict):
            return self._raw_json(msg, transform)
        if isinstance(msg, (bytes, bytearray, str)) and msg[:1] in (b"{", "{"):
            return self._raw_rest_text(msg, transform)
        with self._lock:
            if self._wire is None:
                self._wire = wire.WireStage(self._pick_device())
            st = self._wire
            host = st.host_bytes(msg)
            info = wire.parse_seldon_message(host.numpy())
            X = st.decode_matrix(host, info)
            if shard is not None:
                _, flag = shard[0].score_shard(X, int(shard[1]))
            else:
                _, flag = self.score_device(X)
            if transform:
                self.prediction_meta = flag.cpu().numpy().astype(int)
                off, ln = info["data"]
                resp = wire.build_message(st, info["kind"], None, None, puid=info["puid"],
                                          tags_dev={"outlier-predictions": flag}, metrics=self.metrics(),
                                          data_raw=host.numpy()[off:off + ln])
            else:
                resp = wire.build_message(st, info["kind"], [], flag, puid=info["puid"], tags=self.tags(),
                                          metrics=self.metrics())
        if hasattr(msg, "SerializeToString"):
            return type(msg).FromString(resp.tobytes())
        return resp

    def _raw_rest_text(self, body, transform):
        """REST: the body as TEXT (bytes), before any json.loads - `{"data": {"names": [...], "ndarray": [[...]]}}` or
        `{"data": {"tensor": {"shape": [...], "values": [...]}}}` (utils.py:472-525).  The numbers are converted on the
        device in both directions (decimal text -> binary64 -> fp32, csrc/jsoncodec.cu; flags -> text,
        csrc/jsonencode.cu); only the rest of the body - names, meta - goes through the json module.  Returns the
        response body as bytes, laid out as json.dumps of construct_response_json's dict (utils.py:306-407).  In
        transformer mode the request's array text is echoed as it was sent (the reference re-serialises the same
        numbers)."""
        import json

        import torch

        from . import codec

        if isinstance(body, str):
            body = body.encode("utf-8")
        body = bytes(body)
        with self._lock:
            dev = self._pick_device()
            kind = "tensor" if codec._VALUES_KEY.search(body) and b'"tensor"' in body else "ndarray"
            key = "values" if kind == "tensor" else "ndarray"
            start, end, ndim = codec.json_ndarray_span(body, key)
            rest = json.loads(body[:start] + b"null" + body[end:])
            d = rest.get("data") or {}
            X = codec.decode_json_ndarray(body, dev, key)
            if kind == "tensor":
                shape = [int(v) for v in d["tensor"]["shape"]]
                if len(shape) != 2 or shape[0] * shape[1] != X.numel():
                    raise SeldonB200Error("tensor payload: shape %s does not match %d values" % (shape, X.numel()))
                X = X.view(shape[0], shape[1])
            if X.dim() != 2:
                raise SeldonB200Error("expected a 2-D batch, got shape %s" % (tuple(X.shape),))
            _, flag = self.score_device(X.contiguous())
            holes = {}

            def hole(text):
                k = "@@sb200:%d@@" % len(holes)
                holes[k] = text
                return k

            if transform:
                self.prediction_meta = flag.cpu().numpy().astype(int)
                arr = hole(body[start:end])
                names = d.get("names", [])
            else:
                arr = hole(codec.encode_json_ndarray(flag))
                names = []
            data = {kind: {"values": arr, "shape": list(d["tensor"]["shape"]) if transform else [int(flag.numel())]}
                    if kind == "tensor" else arr, "names": names}
            resp = {"data": data, "meta": {}}
            if transform:
                resp["meta"]["tags"] = {"outlier-predictions": hole(codec.encode_json_ndarray(flag))}
            elif self.tags():
                resp["meta"]["tags"] = self.tags()
            resp["meta"]["metrics"] = self.metrics()
            puid = (rest.get("meta") or {}).get("puid")
            if puid:
                resp["meta"]["puid"] = puid
            out = json.dumps(resp).encode("utf-8")
            for k, text in holes.items():
                out = out.replace(b'"' + k.encode() + b'"', text, 1)
            return out

    def _raw_json(self, request, transform):
        """REST: the body has already been parsed by json.loads (seldon_methods.py:53-57 -> construct_response_json,
        utils.py:306-407): numeric data.tensor / data.ndarray only."""
        d = request.get("data") or {}
        if "tensor" in d:
            X = np.array(d["tensor"]["values"], dtype=np.float64).reshape(d["tensor"]["shape"])
        elif "ndarray" in d:
            X = np.array(d["ndarray"], dtype=np.float64)
        else:
            raise SeldonB200Error("the request carries no data.tensor / data.ndarray payload")
        names = d.get("names", [])
        out = self.transform_input(X, names) if transform else self.predict(X, names)
        out = np.asarray(out)
        data = {}
        if "tensor" in d:
            data["tensor"] = {"values": out.ravel().tolist(), "shape": out.shape}
        else:
            data["ndarray"] = out.tolist()
        data["names"] = list(names) if transform else []
        resp = {"data": data, "meta": {}}
        tags = self.tags()
        if tags:
            resp["meta"]["tags"] = tags
        resp["meta"]["metrics"] = self.metrics()
        puid = (request.get("meta") or {}).get("puid")
        if puid:
            resp["meta"]["puid"] = puid
        return resp

    # ------------------------------------------------------------------ public attributes of the reference
    @property
    def n(self):
        if self._h is None:
            return self._restore["n"] if self._restore else 0
        if getattr(self, "_cached_state", None) is not None:
            return self._cached_state["n"]
        # the row count is mirrored on the host: no need to bring mean, C and the clip bounds back for it
        import torch

        n, hs, hc = ctypes.c_int64(), ctypes.c_int32(), ctypes.c_int32()
        with torch.cuda.device(self._device):
            _cabi.check(_cabi.lib().sb_maha_get_state(self._h, None, None, None, None, ctypes.byref(n), ctypes.byref(hs),
                                                      ctypes.byref(hc), torch.cuda.current_stream().cuda_stream))
        return n.value

    @property
    def mean(self):
        st = self._state()
        return st["mean"] if st["has_state"] else 0  # L35: scalar 0 until the first batch

    @property
    def C(self):
        st = self._state()
        return st["C"] if st["has_state"] else 0

    @property
    def clip(self):
        st = self._state()
        return [st["clip_lo"], st["clip_hi"]] if st["has_clip"] else None

    @property
    def score(self):
        if self._score_t is None:
            raise AttributeError("score")
        return self._score_t.double().cpu().numpy()

    @property
    def prediction(self):
        if self._flag_t is None:
            raise AttributeError("prediction")
        return self._flag_t.cpu().numpy().astype(int)

    # ------------------------------------------------------------------ device path
    def set_option(self, option, value):
        """SB_OPT_* knob of include/sb200.h (applied when the handle exists)."""
        self._options[option] = value
        if self._h is not None:
            _cabi.check(_cabi.lib().sb_maha_set_option(self._h, option, value))

    def fuse_standardise(self, transformer):
        """Score every batch as if `transformer` (a StandardiseB200, or None to switch off) had been applied to it
        first, without the separate pass over the batch: the graph `standardise -> outlier detector` of SURVEY 8
        config 5 in one component.  Results equal those of the two components in sequence (same arithmetic per value;
        scores within the detector's run-to-run spread of a few 1e-7, equal flags); only the
        tensor-core shapes support it (d = 64, 128, 256 and n_components <= 16)."""
        with self._lock:
            self._fused = transformer
            if self._h is not None:
                self._apply_fused()

    def _apply_fused(self):
        t = getattr(self, "_fused", None)
        if t is None:
            _cabi.check(_cabi.lib().sb_maha_set_input_transform(self._h, None, None))
            self._fused_params = None
        else:
            mu, inv = t._params(self._pick_device())
            self._fused_params = (mu, inv)  # keep the device vectors alive while the handle points at them
            _cabi.check(_cabi.lib().sb_maha_set_input_transform(self._h, mu.data_ptr(), inv.data_ptr()))

    def score_device(self, X):
        """One request batch resident on the GPU: X is a CUDA float32 tensor (rows, d), row-major.

        Returns (score float32[rows], flag int32[rows]) CUDA tensors; no host synchronisation."""
        import torch

        with self._lock:
            if not (isinstance(X, torch.Tensor) and X.is_cuda and X.dtype == torch.float32 and X.dim() == 2):
                raise SeldonB200Error("score_device expects a 2-D CUDA float32 tensor")
            if X.stride(1) != 1:
                X = X.contiguous()
            rows, d = X.shape
            if rows < 1:
                raise SeldonB200Error("empty batch")
            self._ensure_handle(d, X.device)
            L = _cabi.lib()
            ws = self._workspace(rows)
            score = torch.empty(rows, dtype=torch.float32, device=X.device)
            flag = torch.empty(rows, dtype=torch.int32, device=X.device)
            with torch.cuda.device(X.device):
                st = torch.cuda.current_stream().cuda_stream
                rc = L.sb_maha_batch(self._h, X.data_ptr(), rows, X.stride(0), score.data_ptr(), flag.data_ptr(),
                                     ws.data_ptr(), ws.numel(), st)
            if rc != 0:
                L.sb_maha_abort_batch(self._h)
            _cabi.check(rc)
            self._score_t, self._flag_t = score, flag
            self._cached_state = None
            return score, flag

    # two-phase form for row-sharded batches (seldon_core_b200/dist.py)
    def pass1_device(self, X, stats_out):
        import torch

        rows, d = X.shape
        self._ensure_handle(d, X.device)
        ws = self._workspace(rows)
        with torch.cuda.device(X.device):
            _cabi.check(_cabi.lib().sb_maha_pass1_stats(self._h, X.data_ptr(), rows, X.stride(0), ws.data_ptr(),
                                                        ws.numel(), stats_out.data_ptr(),
                                                        torch.cuda.current_stream().cuda_stream))

    def finish_device(self, X, all_stats, nranks, rank, total_rows, row0):
        import torch

        rows = X.shape[0]
        L = _cabi.lib()
        ws = self._workspace(rows)
        score = torch.empty(rows, dtype=torch.float32, device=X.device)
        flag = torch.empty(rows, dtype=torch.int32, device=X.device)
        with torch.cuda.device(X.device):
            st = torch.cuda.current_stream().cuda_stream
            try:
                _cabi.check(L.sb_maha_finalize(self._h, all_stats.data_ptr(), nranks, rank, total_rows, row0,
                                               ws.data_ptr(), ws.numel(), st))
                _cabi.check(L.sb_maha_pass2_score(self._h, X.data_ptr(), rows, X.stride(0), score.data_ptr(),
                                                  flag.data_ptr(), ws.data_ptr(), ws.numel(), st))
                _cabi.check(L.sb_maha_commit(self._h, st))
            except Exception:
                L.sb_maha_abort_batch(self._h)
                raise
        self._score_t, self._flag_t = score, flag
        self._cached_state = None
        return score, flag

    def stats_len(self, d):
        return 1 + d + d * d

    # ------------------------------------------------------------------ internals
    def _get_preds(self, X):
        import torch

        with self._lock:
            dev = self._pick_device()
            if isinstance(X, torch.Tensor):
                Xd = X.to(device=dev, dtype=torch.float32)
            else:
                Xn = np.asarray(X)
                if Xn.ndim != 2:
                    raise SeldonB200Error("expected a 2-D batch, got shape %s" % (Xn.shape,))
                if Xn.dtype == np.float64:
                    # the request payload is float64 (utils.py:161-176): ship the 8-byte values and narrow on the device
                    raw = torch.from_numpy(np.ascontiguousarray(Xn)).to(dev, non_blocking=True)
                    Xd = torch.empty(Xn.shape, dtype=torch.float32, device=dev)
                    with torch.cuda.device(dev):
                        _cabi.check(_cabi.lib().sb_unpack_tensor_f64(raw.data_ptr(), 8 * Xn.size, 0, Xn.shape[0], Xn.shape[1], None,
                                                                     None, Xd.data_ptr(), Xn.shape[1],
                                                                     torch.cuda.current_stream().cuda_stream))
                else:
                    Xd = torch.from_numpy(np.ascontiguousarray(Xn, dtype=np.float32)).to(dev, non_blocking=True)
            _, flag = self.score_device(Xd)
            return flag.cpu().numpy().astype(int)

    def _pick_device(self):
        import torch

        if not torch.cuda.is_available():
            raise SeldonB200Error("no CUDA device: seldon_core_b200 has no CPU fallback", status_code=500)
        if self._device is None:
            self._device = torch.device("cuda", torch.cuda.current_device())
        return torch.device(self._device)

    def _ensure_handle(self, d, device):
        if self._h is not None:
            if d != self._d:
                raise SeldonB200Error("feature count changed from %d to %d" % (self._d, d))
            return
        import torch

        self._device = device
        L = _cabi.lib()
        h = ctypes.c_void_p()
        with torch.cuda.device(device):
            _cabi.check(L.sb_maha_create(ctypes.byref(h), d, int(self.n_components), float(self.threshold),
                                         float(self.n_stdev), int(self.start_clip), int(self.max_n)))
        self._h, self._d = h, d
        for opt, val in self._options.items():
            _cabi.check(L.sb_maha_set_option(self._h, opt, val))
        if getattr(self, "_fused", None) is not None:
            self._apply_fused()
        self._cached_state = None
        if self._restore is not None:
            self._push_state(self._restore)
            self._restore = None

    def _workspace(self, rows):
        import torch

        need = ctypes.c_uint64()
        _cabi.check(_cabi.lib().sb_maha_workspace_bytes(self._h, max(int(rows), 1), ctypes.byref(need)))
        if self._ws is None or self._ws.numel() < need.value:
            self._ws = torch.empty(int(need.value), dtype=torch.uint8, device=self._device)
        return self._ws

    def _state(self):
        if self._h is None:
            if self._restore is not None:
                return self._restore
            return {"n": 0, "has_state": 0, "has_clip": 0, "mean": None, "C": None, "clip_lo": None, "clip_hi": None}
        if getattr(self, "_cached_state", None) is not None:
            return self._cached_state
        import torch

        d = self._d
        mean, C, lo, hi = np.empty(d), np.empty((d, d)), np.empty(d), np.empty(d)
        n, hs, hc = ctypes.c_int64(), ctypes.c_int32(), ctypes.c_int32()
        with torch.cuda.device(self._device):
            _cabi.check(_cabi.lib().sb_maha_get_state(self._h, mean.ctypes.data, C.ctypes.data, lo.ctypes.data,
                                                      hi.ctypes.data, ctypes.byref(n), ctypes.byref(hs),
                                                      ctypes.byref(hc), torch.cuda.current_stream().cuda_stream))
        self._cached_state = {"n": n.value, "has_state": hs.value, "has_clip": hc.value, "mean": mean, "C": C,
                              "clip_lo": lo, "clip_hi": hi}
        return self._cached_state

    def _push_state(self, st):
        import torch

        p = lambda a: None if a is None else np.ascontiguousarray(a, dtype=np.float64).ctypes.data
        keep = [np.ascontiguousarray(st[k], dtype=np.float64) if st[k] is not None else None
                for k in ("mean", "C", "clip_lo", "clip_hi")]
        with torch.cuda.device(self._device):
            _cabi.check(_cabi.lib().sb_maha_set_state(
                self._h, *[None if a is None else a.ctypes.data for a in keep], int(st["n"]), int(st["has_state"]),
                int(st["has_clip"]), torch.cuda.current_stream().cuda_stream))
        self._cached_state = None

    def _last_sums(self):
        import torch

        if self._h is None or self._flag_t is None:
            raise AttributeError("prediction")
        nf, ss, gr = ctypes.c_int64(), ctypes.c_double(), ctypes.c_int64()
        with torch.cuda.device(self._device):
            _cabi.check(_cabi.lib().sb_maha_last_batch_sums(self._h, ctypes.byref(nf), ctypes.byref(ss), ctypes.byref(gr),
                                                            torch.cuda.current_stream().cuda_stream))
        self.gate_row = gr.value
        return self._flag_t.numel(), nf.value, ss.value

    def eigvecs(self):
        """Top-k eigenvectors used by the last batch, (d, k) - diagnostics and subspace parity tests."""
        import torch

        k = min(int(self.n_components), self._d)
        V = np.empty((self._d, k))
        with torch.cuda.device(self._device):
            _cabi.check(_cabi.lib().sb_maha_get_eigvecs(self._h, V.ctypes.data, torch.cuda.current_stream().cuda_stream))
        return V

    # ------------------------------------------------------------------ persistence (python/seldon_core/persistence.py:84)
    def __getstate__(self):
        with self._lock:
            st = dict(self._state())
            out = {k: v for k, v in self.__dict__.items() if not k.startswith("_")}
            out["_saved"] = st
            out["_saved_d"] = self._d
            out["_saved_options"] = dict(self._options)
            t = getattr(self, "_fused", None)
            out["_saved_fused"] = None if t is None else (t.mu, t.sigma)
            return out

    def __setstate__(self, state):
        saved = state.pop("_saved")
        d = state.pop("_saved_d")
        opts = state.pop("_saved_options", {})
        fused = state.pop("_saved_fused", None)
        self.__dict__.update(state)
        self._device = None
        self._h = None
        self._d = None
        self._ws = None
        self._lock = threading.RLock()
        self._score_t = None
        self._flag_t = None
        self._options = opts
        self._wire = None
        self._restore = saved if d is not None else None
        self._restore_d = d
        self._fused = None
        if fused is not None:
            from .transformer import StandardiseB200

            self._fused = StandardiseB200(fused[0], fused[1])

    def __del__(self):
        try:
            if self._h is not None:
                _cabi.lib().sb_maha_destroy(self._h)
                self._h = None
        except Exception:
            pass
