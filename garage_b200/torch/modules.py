"""Parameter container for garage's Gaussian MLP (GaussianMLPModule,
torch/modules/gaussian_mlp_module.py:195-305 over MultiHeadedMLPModule,
torch/modules/multi_headed_mlp_module.py:49-151).

The `nn.Module` tree reproduces the reference's attribute names so that
`state_dict()` keys and shapes are interchangeable with the reference's
checkpoints and `agent_update` dicts:
    <prefix>._init_std                                   [1]
    <prefix>._mean_module._layers.{0,1}.linear.{weight,bias}
    <prefix>._mean_module._output_layers.0.linear.{weight,bias}
    <prefix>.min_std_param / max_std_param               buffers (if given)
All parameters are VIEWS into one flat float32 CUDA vector laid out as libb2g
expects ([log_std | W1 | b1 | W2 | b2 | W3 | b3]); the kernels read and Adam
updates that vector in place, `load_state_dict` writes through the views.
"""
import math

import torch
from torch import nn

from garage_b200 import ops

_INF = float('inf')


class _NonLinearity(nn.Module):
    """Parameter-free marker so the Sequential has the reference's layout."""

    def __init__(self, name):
        super().__init__()
        self.name = name


class _MeanModule(nn.Module):

    def __init__(self, input_dim, output_dim, hidden_sizes, hidden_w_init, hidden_b_init,
                 output_w_init, output_b_init):
        super().__init__()
        self._layers = nn.ModuleList()
        prev = input_dim
        for size in hidden_sizes:
            seq = nn.Sequential()
            lin = nn.Linear(prev, size)  # same RNG draws as the reference, then re-initialised
            hidden_w_init(lin.weight)
            hidden_b_init(lin.bias)
            seq.add_module('linear', lin)
            seq.add_module('non_linearity', _NonLinearity('tanh'))
            self._layers.append(seq)
            prev = size
        self._output_layers = nn.ModuleList()
        seq = nn.Sequential()
        lin = nn.Linear(prev, output_dim)
        output_w_init(lin.weight)
        output_b_init(lin.bias)
        seq.add_module('linear', lin)
        self._output_layers.append(seq)


class GaussianMLPModule(nn.Module):
    """Holds the parameters; the arithmetic lives in libb2g."""

    def __init__(self, input_dim, output_dim, hidden_sizes=(32, 32), hidden_nonlinearity=torch.tanh,
                 hidden_w_init=nn.init.xavier_uniform_, hidden_b_init=nn.init.zeros_,
                 output_nonlinearity=None, output_w_init=nn.init.xavier_uniform_,
                 output_b_init=nn.init.zeros_, learn_std=True, init_std=1.0, min_std=1e-6,
                 max_std=None, std_parameterization='exp', layer_normalization=False, device=None):
        super().__init__()
        if std_parameterization not in ('exp', 'softplus'):
            raise NotImplementedError
        # what the CUDA path is built for; anything else fails loudly (no fallback)
        if hidden_nonlinearity not in (torch.tanh, ) and not isinstance(hidden_nonlinearity, nn.Tanh):
            raise NotImplementedError('garage_b200 kernels implement tanh hidden layers only')
        if output_nonlinearity is not None or layer_normalization or std_parameterization != 'exp' \
                or not learn_std:
            raise NotImplementedError('garage_b200 kernels implement the default GaussianMLPModule '
                                      '(linear output, exp std, learned scalar std, no layer norm)')
        hidden_sizes = tuple(int(h) for h in hidden_sizes)
        if len(hidden_sizes) != 2 or hidden_sizes[0] != hidden_sizes[1]:
            raise NotImplementedError('garage_b200 kernels need hidden_sizes=(H, H); got {}'.format(hidden_sizes))
        self._input_dim, self._action_dim, self._hidden = int(input_dim), int(output_dim), hidden_sizes[0]
        self.device = device or torch.device('cuda', torch.cuda.current_device())
        self.shape = ops.MlpShape(self._input_dim, self._hidden, self._action_dim, self.device)

        # initialise on the CPU with the reference's sequence of RNG draws
        self._init_std = nn.Parameter(torch.Tensor([init_std]).log())
        self._mean_module = _MeanModule(input_dim, output_dim, hidden_sizes, hidden_w_init,
                                        hidden_b_init, output_w_init, output_b_init)
        self.min_log_std, self.max_log_std = -_INF, _INF
        if min_std is not None:
            self.register_buffer('min_std_param', torch.Tensor([min_std]).log())
        if max_std is not None:
            self.register_buffer('max_std_param', torch.Tensor([max_std]).log())
        # gaussian_mlp_module.py:171: `if self._min_std_param or self._max_std_param` is tensor
        # truthiness, so a bound whose log is exactly 0 does not switch the clamp on (Q6)
        lo = None if min_std is None else float(self.min_std_param.item())
        hi = None if max_std is None else float(self.max_std_param.item())
        if (lo is not None and lo != 0.0) or (hi is not None and hi != 0.0):
            self.min_log_std = -_INF if lo is None else lo
            self.max_log_std = _INF if hi is None else hi
        self._flatten_to_device()

    def _ordered(self):
        m = self._mean_module
        return [('_init_std', self), ('weight', m._layers[0].linear), ('bias', m._layers[0].linear),
                ('weight', m._layers[1].linear), ('bias', m._layers[1].linear),
                ('weight', m._output_layers[0].linear), ('bias', m._output_layers[0].linear)]

    def _flatten_to_device(self):
        flat = torch.empty(self.shape.P, dtype=torch.float32, device=self.device)
        off = 0
        for name, owner in self._ordered():
            p = getattr(owner, name)
            n = p.numel()
            flat[off:off + n].copy_(p.detach().reshape(-1))
            view = nn.Parameter(flat[off:off + n].view(p.shape), requires_grad=True)
            owner._parameters[name] = view
            off += n
        assert off == self.shape.P
        self.flat = flat
        for name in ('min_std_param', 'max_std_param'):
            if name in self._buffers and self._buffers[name] is not None:
                self._buffers[name] = self._buffers[name].to(self.device)

    # the flat vector is the single source of truth: rebuild views after unpickling
    def __getstate__(self):
        state = self.__dict__.copy()
        state['flat'] = None
        state['shape'] = None
        state['_cpu_params'] = {k: v.detach().cpu() for k, v in self.state_dict().items()}
        return state

    def __setstate__(self, state):
        cpu = state.pop('_cpu_params')
        self.__dict__.update(state)
        self.device = torch.device('cuda', torch.cuda.current_device())
        self.shape = ops.MlpShape(self._input_dim, self._hidden, self._action_dim, self.device)
        for name, owner in self._ordered():
            owner._parameters[name] = nn.Parameter(owner._parameters[name].detach().cpu())
        self._flatten_to_device()
        self.load_state_dict(cpu)

    def log_std_value(self):
        """Clamped scalar log-std as a python float (one tiny D2H)."""
        return min(max(float(self.flat[0].item()), self.min_log_std), self.max_log_std)

    def mean(self, obs):
        """Batched MLP forward on the GPU: obs [n,O] float32 CUDA -> mean [n,A]."""
        out = torch.empty(obs.shape[0], self._action_dim, dtype=torch.float32, device=self.device)
        if obs.shape[0]:
            ops.mlp_eval(self.shape, self.flat, obs, mean_out=out)
        return out
