"""Drop-ins for the reference's multi-armed-bandit routers, decision-for-decision identical under a seed.

Mirror ``components/routers/epsilon-greedy/EpsilonGreedy.py`` (L35-136) and
``components/routers/thompson-sampling/ThompsonSampling.py`` (L35-115): same constructor parameters, ``route`` returns
a Python ``int`` (the wrapper rejects anything else, ``python/seldon_core/seldon_methods.py:231``), ``send_feedback``
accepts both argument orders found in the reference tree:

* component style  ``send_feedback(features, feature_names, routing, reward, truth)``  (EpsilonGreedy.py:108)
* wrapper style    ``send_feedback(features, feature_names, reward, truth, routing=r)`` (user_model.py:308-313)

The bandit state and both MT19937 streams live on the GPU (csrc/routers.cu); ``run_batched`` plays T interleaved
route/feedback calls in one launch.  A falsy seed leaves the reference on whatever the process-global generators
hold; here a seed is then drawn from os.urandom (decisions are not reproducible in either implementation).
"""
import ctypes
import os
import threading

import numpy as np

from . import _cabi
from .errors import SeldonB200Error


class _RouterBase(object):
    _kind = None

    def _init_common(self, n_branches, epsilon, best_branch, verbose, seed, history, branch_names, device):
        try:
            n_branches = int(n_branches)  # TypeError for None, like the reference
            assert n_branches > 0
        except (TypeError, ValueError, AssertionError):
            raise
        self.verbose = verbose
        self.n_branches = n_branches
        self.history = history
        self.seed = seed
        if self.history:
            self.branch_history = []
            self.value_history = []
        if branch_names is not None:
            self.branch_names = branch_names.split(":")
        self._eff_seed = int(seed) if seed else (int.from_bytes(os.urandom(4), "little") | 1)
        self._epsilon = float(epsilon)
        self._init_best = int(best_branch) if best_branch else 0
        self._device = device
        self._h = None
        self._lock = threading.RLock()
        self._restore = None
        self._pending = None   # (routing, rows, reward) of a send_feedback not yet launched: it rides with the next route

    # -- device handle
    def _handle(self):
        if self._h is None:
            import torch

            if not torch.cuda.is_available():
                raise SeldonB200Error("no CUDA device: seldon_core_b200 has no CPU fallback", status_code=500)
            if self._device is None:
                self._device = torch.device("cuda", torch.cuda.current_device())
            self._device = torch.device(self._device)
            if self._device.index is None:
                self._device = torch.device("cuda", torch.cuda.current_device())
            h = ctypes.c_void_p()
            with torch.cuda.device(self._device):
                _cabi.check(_cabi.lib().sb_router_create(ctypes.byref(h), self._kind, self.n_branches, self._epsilon,
                                                         self._eff_seed, self._init_best))
            self._h = h
            if self._restore is not None:
                self._set_state(self._restore)
                self._restore = None
        return self._h

    def _stream(self):
        import torch

        return torch.cuda.current_stream(self._device).cuda_stream

    def _flush(self):
        """Launch a feedback that was waiting for the next route (called before anything else touches the state)."""
        import torch

        if self._pending is not None:
            routing, rows, reward = self._pending
            self._pending = None
            with torch.cuda.device(self._device):
                _cabi.check(_cabi.lib().sb_router_feedback(self._handle(), routing, rows, reward, self._stream()))

    def _get_state(self):
        import torch

        h = self._handle()
        self._flush()
        L = _cabi.lib()
        st = np.zeros(L.sb_router_state_len(h), dtype=np.int64)
        with torch.cuda.device(self._device):
            _cabi.check(L.sb_router_get_state(h, st.ctypes.data, self._stream()))
        return st

    def _set_state(self, st):
        import torch

        st = np.ascontiguousarray(st, dtype=np.int64)
        with torch.cuda.device(self._device):
            _cabi.check(_cabi.lib().sb_router_set_state(self._h, st.ctypes.data, self._stream()))

    @property
    def branch_success(self):
        return self._get_state()[: self.n_branches].tolist()

    @property
    def branch_tries(self):
        return self._get_state()[self.n_branches: 2 * self.n_branches].tolist()

    # -- reference API
    def route(self, features, feature_names):
        import torch

        with self._lock:
            h = self._handle()
            out = ctypes.c_int32()
            L = _cabi.lib()
            pend, self._pending = self._pending, None
            if torch.cuda.current_device() != self._device.index:
                torch.cuda.set_device(self._device)
            st = self._stream()
            if pend is None:
                rc = L.sb_router_route(h, ctypes.byref(out), st)
            else:  # the previous request's feedback and this route in one launch
                rc = L.sb_router_feedback_route(h, pend[0], pend[1], pend[2], ctypes.byref(out), st)
            if rc:
                _cabi.check(rc)
            branch = int(out.value)
            if self.history:  # EpsilonGreedy.py:99-101, ThompsonSampling.py:88-90
                self.branch_history.append(branch)
                self.value_history.append(list(self.branch_values))
            return branch

    def counters(self):
        """{guard_hits, fallbacks}: comparisons of the legacy gamma sampler decided again with correctly rounded
        logarithms, and beta draws taken draw by draw outside the look-ahead window (include/sb200.h)."""
        import torch

        self._handle()
        self._flush()
        out = (ctypes.c_int64 * 2)()
        with torch.cuda.device(self._device):
            _cabi.check(_cabi.lib().sb_router_counters(self._handle(), out, self._stream()))
        return {"guard_hits": int(out[0]), "fallbacks": int(out[1])}

    def send_feedback(self, features, feature_names, *args, routing=None, **kwargs):
        if routing is None and len(args) == 3:  # component style: (routing, reward, truth)
            routing, reward, truth = args
        elif len(args) >= 1:  # wrapper style: (reward, truth) + routing keyword
            reward = args[0]
        else:
            reward = kwargs.get("reward")
        if routing is None or reward is None:
            raise SeldonB200Error("send_feedback needs a routing and a reward")
        rows = int(features.shape[0])  # n_success_failures, EpsilonGreedy.py:132-136
        import torch

        with self._lock:
            h = self._handle()
            routing = int(routing)
            if not 0 <= routing < self.n_branches:
                raise SeldonB200Error("send_feedback: routing %d outside 0..%d" % (routing, self.n_branches - 1))
            if rows < 1:
                raise SeldonB200Error("send_feedback: rows must be >= 1 (reference divides by tries)")
            self._flush()
            # not launched yet: the next route() applies it in its own launch (sb_router_feedback_route); anything else
            # that reads or changes the state flushes it first
            self._pending = (routing, rows, float(reward))
            del h

    def n_success_failures(self, features, reward):
        n_predictions = features.shape[0]
        n_success = int(reward * n_predictions)
        return n_success, n_predictions - n_success

    def run_batched(self, T, reward, reward_mode=_cabi.SB_REWARD_PER_CALL, arm_p=None, rows_per_call=None):
        """T interleaved route()/send_feedback() calls in one launch.  `reward` (float32) and the optional
        `arm_p` / `rows_per_call` may be numpy arrays or CUDA tensors.  Returns the decisions as a CUDA int32 tensor."""
        import torch

        with self._lock:
            h = self._handle()
            self._flush()
            dev = self._device

            def to_dev(a, dt):
                if a is None:
                    return None
                if isinstance(a, torch.Tensor):
                    return a.to(device=dev, dtype=dt).contiguous()
                return torch.from_numpy(np.ascontiguousarray(a)).to(device=dev, dtype=dt)

            rew = to_dev(reward, torch.float32)
            ap = to_dev(arm_p, torch.float32)
            rw = to_dev(rows_per_call, torch.int64)
            dec = torch.empty(int(T), dtype=torch.int32, device=dev)
            ptr = lambda t: None if t is None else t.data_ptr()
            with torch.cuda.device(dev):
                _cabi.check(_cabi.lib().sb_router_run(h, int(T), ptr(rw), ptr(rew), ptr(ap), int(reward_mode),
                                                      dec.data_ptr(), self._stream()))
            return dec

    # -- persistence (python/seldon_core/persistence.py:84): bandit state and both generator positions
    def __getstate__(self):
        with self._lock:
            out = {k: v for k, v in self.__dict__.items() if k not in ("_h", "_lock", "_device", "_restore", "_pending")}
            out["_saved"] = self._get_state() if self._h is not None else self._restore
            out["_pending_saved"] = None
            return out

    def __setstate__(self, state):
        saved = state.pop("_saved")
        state.pop("_pending_saved", None)
        self.__dict__.update(state)
        self._pending = None
        self._h = None
        self._device = None
        self._lock = threading.RLock()
        self._restore = saved

    def __del__(self):
        try:
            if self._h is not None:
                _cabi.lib().sb_router_destroy(self._h)
                self._h = None
        except Exception:
            pass


class EpsilonGreedyB200(_RouterBase):
    _kind = _cabi.SB_ROUTER_EPSILON_GREEDY

    def __init__(self, n_branches=None, epsilon=0.1, best_branch=None, verbose=False, seed=None, history=False,
                 branch_names=None, device=None):
        self.epsilon = epsilon
        self._init_common(n_branches, epsilon, best_branch, verbose, seed, history, branch_names, device)

    @property
    def best_branch(self):
        return int(self._get_state()[2 * self.n_branches])

    @property
    def branch_values(self):
        st = self._get_state()
        s, t = st[: self.n_branches], st[self.n_branches: 2 * self.n_branches]
        return [int(a) / int(b) if b else 0 for a, b in zip(s, t)]


class ThompsonSamplingB200(_RouterBase):
    _kind = _cabi.SB_ROUTER_THOMPSON

    def __init__(self, n_branches=None, verbose=False, seed=None, history=False, branch_names=None, device=None):
        self._init_common(n_branches, 0.0, None, verbose, seed, history, branch_names, device)

    @property
    def branch_values(self):
        """The beta samples of the last route (ThompsonSampling.py:81-82); zeros before the first one (L65)."""
        import torch

        v = np.zeros(self.n_branches)
        self._handle()
        with torch.cuda.device(self._device):
            _cabi.check(_cabi.lib().sb_router_last_values(self._handle(), v.ctypes.data, self._stream()))
        return v.tolist()

    @property
    def models_beta_params(self):
        st = self._get_state()
        s, t = st[: self.n_branches], st[self.n_branches: 2 * self.n_branches]
        return [[1 + int(a), 1 + int(b - a)] for a, b in zip(s, t)]
