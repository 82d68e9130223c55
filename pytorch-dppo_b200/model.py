"""Actor-critic model and the shared buffers of the reference's model.py, with
storage laid out for the B200 kernels.

``Model`` keeps the reference's interface (model.py:8-45): attributes
``p_fc1, p_fc2, v_fc1, v_fc2, mu, v`` with ``.weight (out,in)`` / ``.bias``,
``log_std (1,A)``, ``forward(x) -> (mu, sigma_sq = exp(log_std), v)`` and the
13 ``state_dict`` keys in the reference's order - but all 13 tensors are VIEWS
into one flat fp32 bucket (``model.flat``) in the layout of
``dppo_param_layout`` (include/dppo_b200.h), which is what K2/K3 read and
write.  Hidden widths default to the reference's hard-coded 100/100 and
500/100 (model.py:11-18) and can be passed explicitly for the 64x64 / 256x256
benchmark configs.

``forward`` is plain torch on the views (it serves rollouts and evaluation,
which are outside the update path); the update itself never goes through
autograd - see b200_engine.UpdateEngine."""
import torch
import torch.nn as nn
import torch.nn.functional as F

import b200_native as nv


def default_device():
    return torch.device("cuda", torch.cuda.current_device()) if torch.cuda.is_available() else torch.device("cpu")


class _Affine(nn.Module):
    """nn.Linear-shaped holder whose weight and bias are views into the flat bucket."""

    def __init__(self, weight, bias):
        super().__init__()
        self.weight = nn.Parameter(weight)
        self.bias = nn.Parameter(bias)
        self.out_features, self.in_features = weight.shape

    def forward(self, x):
        return F.linear(x, self.weight, self.bias)


class Model(nn.Module):
    def __init__(self, num_inputs, num_outputs, h_size_1=100, h_size_2=100, v_h_size_1=None, v_h_size_2=None,
                 device=None):
        super().__init__()
        v_h_size_1 = 5 * h_size_1 if v_h_size_1 is None else v_h_size_1       # model.py:17
        v_h_size_2 = h_size_2 if v_h_size_2 is None else v_h_size_2
        self.dims = nv.Dims(num_inputs, num_outputs, h_size_1, h_size_2, v_h_size_1, v_h_size_2)
        self.offsets, self.sizes, self.P_pad = nv.param_layout(self.dims)
        self.num_params = sum(self.sizes)
        device = default_device() if device is None else torch.device(device)

        # Same RNG consumption as the reference constructor (model.py:14-22): nn.Linear
        # default init in the order p_fc1, p_fc2, v_fc1, v_fc2, mu, v on the CPU generator.
        init = {}
        for name, (i, o) in (("p_fc1", (num_inputs, h_size_1)), ("p_fc2", (h_size_1, h_size_2)),
                             ("v_fc1", (num_inputs, v_h_size_1)), ("v_fc2", (v_h_size_1, v_h_size_2)),
                             ("mu", (h_size_2, num_outputs)), ("v", (v_h_size_2, 1))):
            init[name + ".weight"] = nn.Linear(i, o).weight.detach()
        flat = torch.zeros(self.P_pad, dtype=torch.float32)                 # biases 0 (model.py:24-27), log_std 0
        shapes = self.dims.shapes()
        for i, name in enumerate(nv.PARAM_NAMES):
            if name in init:
                flat[self.offsets[i]:self.offsets[i] + self.sizes[i]] = init[name].reshape(-1)
        self.flat = flat.to(device)

        def view(name):
            i = nv.PARAM_NAMES.index(name)
            return self.flat[self.offsets[i]:self.offsets[i] + self.sizes[i]].view(shapes[name])

        self.log_std = nn.Parameter(view("log_std"))                         # registered first, as in the reference
        self.p_fc1 = _Affine(view("p_fc1.weight"), view("p_fc1.bias"))
        self.p_fc2 = _Affine(view("p_fc2.weight"), view("p_fc2.bias"))
        self.v_fc1 = _Affine(view("v_fc1.weight"), view("v_fc1.bias"))
        self.v_fc2 = _Affine(view("v_fc2.weight"), view("v_fc2.bias"))
        self.mu = _Affine(view("mu.weight"), view("mu.bias"))
        self.v = _Affine(view("v.weight"), view("v.bias"))
        self.train()                                                         # model.py:33

    def forward(self, inputs):
        x = torch.tanh(self.p_fc1(inputs))
        x = torch.tanh(self.p_fc2(x))
        mu = self.mu(x)
        sigma_sq = torch.exp(self.log_std)            # named sigma_sq, is exp(log_std) (model.py:40)
        x = torch.tanh(self.v_fc1(inputs))
        x = torch.tanh(self.v_fc2(x))
        return mu, sigma_sq, self.v(x)

    def act(self, states, eps=None, std_is_variance=True, mu=None, v=None, act=None, logp=None, states_out=None,
            row0=0, row_stride=1):
        """The rollout's forward for a BATCH of normalised states on the update path's own kernel (no autograd, no
        cuBLAS): mu, v, ``action = mu + std*eps`` and its log-probability (train.py:87-90 / ppo.py:91-98).  Outputs
        that are given are filled in place (row r of the batch -> row ``row0 + r*row_stride``); returns (mu, v, act,
        logp) tensors of the batch when none is given."""
        states = states.reshape(-1, self.dims.O).contiguous()
        N = states.shape[0]
        own = mu is None and v is None and act is None and logp is None
        if own:
            mu = torch.empty(N, self.dims.A, device=self.flat.device)
            act = torch.empty(N, self.dims.A, device=self.flat.device)
            v = torch.empty(N, device=self.flat.device)
            logp = torch.empty(N, device=self.flat.device)
        nv.policy_forward(self.dims, self.flat, states, eps, std_is_variance, mu, v, act, logp, states_out, row0, row_stride)
        return (mu, v, act, logp) if own else None

    @property
    def device(self):
        return self.flat.device

    def share_memory(self):
        """The reference shares the model between forked processes (main.py:52); here
        every GPU process holds a replica that K3 keeps bit-identical, so this is a no-op
        on CUDA."""
        if self.flat.is_cuda:
            return self
        return super().share_memory()

    def _apply(self, fn, recurse=True):
        moved = fn(self.flat)
        if moved.device != self.flat.device or moved.dtype != self.flat.dtype:
            raise RuntimeError("Model lives in one flat bucket: construct it with device=... instead of .to()/.cuda()")
        return self


class Shared_grad_buffers():
    """reference model.py:47-59.  ``grads[name+'_grad']`` are views into one flat
    bucket (same layout as ``Model.flat``), initialised to ONES like the reference
    (model.py:51; padding floats are 0).  K3 reads the bucket as the running sum that
    the workers' gradients are added to and zeroes it after the Adam step
    (chief.py:19)."""

    def __init__(self, model):
        self.flat = torch.zeros_like(model.flat)
        self.grads = {}
        for i, (name, p) in enumerate(model.named_parameters()):
            v = self.flat[model.offsets[i]:model.offsets[i] + model.sizes[i]].view(p.shape)
            v.fill_(1.0)
            self.grads[name + '_grad'] = v

    def add_gradient(self, model):
        flat_grad = getattr(model, "flat_grad", None)
        if flat_grad is not None:                    # gradient produced by the fused kernels
            self.flat += flat_grad
            return
        for name, p in model.named_parameters():     # gradient produced by autograd on the views
            self.grads[name + '_grad'] += p.grad.data

    def reset(self):
        self.flat.zero_()


class Shared_obs_stats():
    """reference model.py:61-80: running per-feature n / mean / M2 / var.  ``observes``
    and ``normalize`` keep the one-row interface; ``observes_batch`` runs a whole
    rollout through K4 with the same prefix semantics (row t normalised with the
    statistics of rows 0..t)."""

    def __init__(self, num_inputs, device=None):
        device = default_device() if device is None else torch.device(device)
        self.n = torch.zeros(num_inputs, device=device)
        self.mean = torch.zeros(num_inputs, device=device)
        self.mean_diff = torch.zeros(num_inputs, device=device)
        self.var = torch.zeros(num_inputs, device=device)
        self._scratch = None

    def _scratch_for(self, T):
        need = nv.welford_scratch_bytes(T, self.n.numel())
        if self._scratch is None or self._scratch.numel() < need:
            self._scratch = torch.empty(need, dtype=torch.uint8, device=self.n.device)
        return self._scratch

    def _shared_for(self, T, world):
        """Buffers of the rank-ordered sharing across workers: this worker's totals, all workers' totals, scratch."""
        O = self.n.numel()
        need = nv.welford_totals_scratch_bytes(T, O)
        sh = getattr(self, "_shared", None)
        if sh is None or sh[1].shape[0] != world or sh[2].numel() < need:
            dev = self.n.device
            sh = self._shared = (torch.zeros(1 + 2 * O, dtype=torch.float64, device=dev),
                                 torch.zeros(world, 1 + 2 * O, dtype=torch.float64, device=dev),
                                 torch.empty(max(need, 16), dtype=torch.uint8, device=dev))
        return sh

    def observes_batch(self, obs, out=None, comm=None):
        """A whole rollout through K4.  With ``comm`` (W > 1 workers, one per GPU) the statistics are the ONE set
        the reference's workers share (main.py:54): the W rollouts of this outer iteration enter in rank order -
        worker r continues from the rollouts of workers 0..r-1 - and every replica ends with the same state."""
        obs = obs.reshape(-1, self.n.numel()).contiguous()
        out = torch.empty_like(obs) if out is None else out
        if comm is not None and comm.world > 1:
            mine, every, scratch = self._shared_for(obs.shape[0], comm.world)
            nv.welford_totals(obs, self.mean, mine, scratch)
            comm.allgather(mine, every)
            nv.welford_shared(obs, out, self.n, self.mean, self.mean_diff, self.var, every, comm.rank,
                              self._scratch_for(obs.shape[0]))
        else:
            nv.welford(obs, out, self.n, self.mean, self.mean_diff, self.var, self._scratch_for(obs.shape[0]))
        return out

    def clone(self):
        """An independent copy of the statistics (the evaluator normalises with its own copy so that the
        replicas of the training statistics stay identical on every rank)."""
        c = Shared_obs_stats(self.n.numel(), device=self.n.device)
        for k in ("n", "mean", "mean_diff", "var"):
            getattr(c, k).copy_(getattr(self, k))
        return c

    def observes(self, obs):
        x = getattr(obs, "data", obs).reshape(1, -1).to(self.n.device, torch.float32)
        self.observes_batch(x)

    def normalize(self, inputs):
        return torch.clamp((inputs - self.mean) / torch.sqrt(self.var), -5., 5.)
