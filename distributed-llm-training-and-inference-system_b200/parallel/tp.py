"""Tensor + sequence parallelism (Megatron-style) building blocks.

* ``ColumnParallelLinear`` / ``RowParallelLinear`` — weights sharded over the TP group.
* With ``sequence_parallel`` the activations outside attention/MLP are sharded on the token dim:
  all-gather before the column-parallel GEMM and reduce-scatter after the row-parallel GEMM
  (SURVEY.md §5.7a).  On CUDA these two pairs are the fused K1 ``ag_gemm`` / K2 ``gemm_rs`` kernels
  (``comms/fused.py``); the NCCL/gloo collectives below are the baseline path and the CPU path.
* ``VocabParallelEmbedding`` and ``vocab_parallel_cross_entropy`` (K15).

Reference: TP exists only as a planner integer (reference llmctl/cli/commands/plan.py:79-80,110-113,136-138);
``sequence_parallel`` only as a preset flag (reference configs/presets/llama-7b-a100x8.toml:36).
"""
from __future__ import annotations


import torch
import torch.distributed as dist
import torch.nn as nn
import torch.nn.functional as F

from ..comms.mesh import get_state
from ..ops import functional as ops


# ------------------------------------------------------------------------------------------------
# collective autograd functions (baseline transport: torch.distributed)
# ------------------------------------------------------------------------------------------------
def _tp():
    st = get_state()
    return st.tp_group, st.tp


class _CopyToTP(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x):
        return x

    @staticmethod
    def backward(ctx, g):
        grp, n = _tp()
        if n > 1:
            g = g.contiguous()
            dist.all_reduce(g, group=grp)
        return g


class _ReduceFromTP(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x):
        grp, n = _tp()
        if n > 1:
            x = x.contiguous()
            dist.all_reduce(x, group=grp)
        return x

    @staticmethod
    def backward(ctx, g):
        return g


def _all_gather_dim0(x: torch.Tensor) -> torch.Tensor:
    grp, n = _tp()
    x = x.contiguous()
    out = torch.empty((x.shape[0] * n, *x.shape[1:]), dtype=x.dtype, device=x.device)
    dist.all_gather_into_tensor(out, x, group=grp)
    return out


def _reduce_scatter_dim0(x: torch.Tensor) -> torch.Tensor:
    grp, n = _tp()
    x = x.contiguous()
    out = torch.empty((x.shape[0] // n, *x.shape[1:]), dtype=x.dtype, device=x.device)
    if dist.get_backend(grp) == "gloo":
        # gloo has no reduce_scatter: all-reduce then slice (CPU plumbing path only)
        dist.all_reduce(x, group=grp)
        r = get_state().tp_rank
        out.copy_(x[r * out.shape[0]:(r + 1) * out.shape[0]])
    else:
        dist.reduce_scatter_tensor(out, x, group=grp)
    return out


class _GatherSeq(torch.autograd.Function):
    """SP: all-gather tokens forward, reduce-scatter gradients backward."""

    @staticmethod
    def forward(ctx, x):
        return _all_gather_dim0(x) if _tp()[1] > 1 else x

    @staticmethod
    def backward(ctx, g):
        return _reduce_scatter_dim0(g) if _tp()[1] > 1 else g


class _ReduceScatterSeq(torch.autograd.Function):
    """SP: reduce-scatter partial sums forward, all-gather gradients backward."""

    @staticmethod
    def forward(ctx, x):
        return _reduce_scatter_dim0(x) if _tp()[1] > 1 else x

    @staticmethod
    def backward(ctx, g):
        return _all_gather_dim0(g) if _tp()[1] > 1 else g


class _ScatterSeq(torch.autograd.Function):
    """Take this rank's token shard (input replicated); backward all-gathers."""

    @staticmethod
    def forward(ctx, x):
        st = get_state()
        if st.tp == 1:
            return x
        n = x.shape[0] // st.tp
        return x[st.tp_rank * n:(st.tp_rank + 1) * n].contiguous()

    @staticmethod
    def backward(ctx, g):
        return _all_gather_dim0(g) if _tp()[1] > 1 else g


copy_to_tp = _CopyToTP.apply
reduce_from_tp = _ReduceFromTP.apply
gather_seq = _GatherSeq.apply
reduce_scatter_seq = _ReduceScatterSeq.apply
scatter_seq = _ScatterSeq.apply


# ------------------------------------------------------------------------------------------------
# layers
# ------------------------------------------------------------------------------------------------
def _init_weight(shape, std, dtype, device, seed_offset: int = 0):
    w = torch.empty(shape, dtype=dtype, device=device)
    nn.init.normal_(w, mean=0.0, std=std)
    return w


class ColumnParallelLinear(nn.Module):
    """y_local = x @ W_local.T, W sharded on the output dim.  Input is replicated over TP, or — with
    sequence parallelism — sharded on tokens and all-gathered first."""

    def __init__(self, in_features: int, out_features: int, *, std: float = 0.02, dtype=torch.bfloat16,
                 device=None, fused_name: str = "", packed: int = 1):
        super().__init__()
        st = get_state()
        assert out_features % (st.tp * packed) == 0, f"{out_features} not divisible by tp={st.tp} x packed={packed}"
        self.in_features, self.out_features = in_features, out_features
        self.out_local = out_features // st.tp
        self.weight = nn.Parameter(_init_weight((self.out_local, in_features), std, dtype, device))
        self.weight.tp_shard_dim = 0
        self.weight.tp_packed = packed      # rows are `packed` equal groups (q|k|v, gate|up), each sharded
        self.fused_name = fused_name

    def forward(self, x: torch.Tensor) -> torch.Tensor:
        st = get_state()
        if st.tp > 1:
            if st.sequence_parallel:
                from ..comms import fused
                if fused.enabled_ag(x):
                    return fused.ag_linear(x, self.weight)
                x = gather_seq(x)
            else:
                x = copy_to_tp(x)
        return ops.linear(x, self.weight)


class RowParallelLinear(nn.Module):
    """y = sum_tp(x_local @ W_local.T), W sharded on the input dim; output all-reduced, or — with
    sequence parallelism — reduce-scattered on tokens."""

    def __init__(self, in_features: int, out_features: int, *, std: float = 0.02, dtype=torch.bfloat16,
                 device=None, fused_name: str = ""):
        super().__init__()
        st = get_state()
        assert in_features % st.tp == 0, f"{in_features} not divisible by tp={st.tp}"
        self.in_features, self.out_features = in_features, out_features
        self.in_local = in_features // st.tp
        self.weight = nn.Parameter(_init_weight((out_features, self.in_local), std, dtype, device))
        self.weight.tp_shard_dim = 1
        self.fused_name = fused_name

    def forward(self, x: torch.Tensor) -> torch.Tensor:
        st = get_state()
        if st.tp > 1 and st.sequence_parallel:
            from ..comms import fused
            if fused.enabled_rs(x, self.weight.shape[0]):
                return fused.linear_rs(x, self.weight)
        return self.finish(ops.linear(x, self.weight))

    def finish(self, y_local: torch.Tensor) -> torch.Tensor:
        """Reduce the per-rank partial product (all-reduce, or reduce-scatter over tokens with SP)."""
        st = get_state()
        if st.tp > 1:
            return reduce_scatter_seq(y_local) if st.sequence_parallel else reduce_from_tp(y_local)
        return y_local


class VocabParallelEmbedding(nn.Module):
    def __init__(self, vocab_size: int, hidden: int, *, std: float = 0.02, dtype=torch.bfloat16, device=None):
        super().__init__()
        st = get_state()
        assert vocab_size % st.tp == 0
        self.vocab_size, self.hidden = vocab_size, hidden
        self.v_local = vocab_size // st.tp
        self.v_start = st.tp_rank * self.v_local
        self.weight = nn.Parameter(_init_weight((self.v_local, hidden), std, dtype, device))
        self.weight.tp_shard_dim = 0

    def forward(self, ids: torch.Tensor) -> torch.Tensor:
        st = get_state()
        if st.tp == 1:
            return F.embedding(ids, self.weight)
        local = ids - self.v_start
        mask = (local < 0) | (local >= self.v_local)
        local = local.masked_fill(mask, 0)
        out = F.embedding(local, self.weight)
        out = out.masked_fill(mask.unsqueeze(-1), 0)
        flat = out.reshape(-1, self.hidden)
        if st.sequence_parallel:
            return reduce_scatter_seq(flat)
        return reduce_from_tp(flat).view(*ids.shape, self.hidden)


class _VocabParallelCE(torch.autograd.Function):
    """Cross-entropy over logits sharded on the vocab dim (K15): local max / sum-exp / target logit,
    two tiny all-reduces, gradient written back in place."""

    @staticmethod
    def forward(ctx, logits, targets, v_start, ignore_index, grad_scale):
        grp, n = _tp()
        lf = logits.float()
        local_max = lf.max(dim=-1).values
        gmax = local_max.clone()
        dist.all_reduce(gmax, op=dist.ReduceOp.MAX, group=grp)
        ex = (lf - gmax.unsqueeze(-1)).exp()
        sumexp = ex.sum(-1)
        v_local = logits.shape[-1]
        t_local = targets - v_start
        own = (t_local >= 0) & (t_local < v_local)
        t_safe = t_local.masked_fill(~own, 0)
        tgt_logit = lf.gather(-1, t_safe.unsqueeze(-1)).squeeze(-1).masked_fill(~own, 0.0)
        packed = torch.stack([sumexp, tgt_logit])
        dist.all_reduce(packed, group=grp)
        sumexp, tgt_logit = packed[0], packed[1]
        valid = targets != ignore_index
        cnt = valid.sum().clamp_min(1).float()
        loss_rows = (gmax + sumexp.log() - tgt_logit).masked_fill(~valid, 0.0)
        loss = loss_rows.sum() / cnt * grad_scale
        soft = ex / sumexp.unsqueeze(-1)
        onehot_rows = torch.nonzero(own & valid).squeeze(-1)
        soft[onehot_rows, t_safe[onehot_rows]] -= 1.0
        soft = soft * (valid.float() / cnt * grad_scale).unsqueeze(-1)
        ctx.save_for_backward(soft.to(logits.dtype))
        return loss

    @staticmethod
    def backward(ctx, dloss):
        (g,) = ctx.saved_tensors
        return g * dloss.to(g.dtype), None, None, None, None


class _VocabParallelCENative(torch.autograd.Function):
    """K15 on the TP path: the two native kernels of csrc/loss/cross_entropy.cu around two tiny all-reduces.
    ``ce_partial`` reads the bf16 shard once (local max, sum-exp, owned target logit per row); the [rows]-sized
    statistics are combined across the TP group (max, then sum); ``ce_finalize`` overwrites the shard in place with
    its gradient.  No fp32 [rows, V/tp] tensor is ever materialised."""

    @staticmethod
    def forward(ctx, logits, targets, v_start, ignore_index, grad_scale):
        from ..ops import native
        grp, n = _tp()
        rows, v_local = logits.shape
        stats = torch.empty(rows, 3, dtype=torch.float32, device=logits.device)
        native.call("llmctl_cross_entropy_partial", logits.data_ptr(), logits.stride(0), targets.data_ptr(),
                    stats.data_ptr(), rows, v_local, int(v_start), native.stream_ptr())
        local_max = stats[:, 0]
        gmax = local_max.clone()
        dist.all_reduce(gmax, op=dist.ReduceOp.MAX, group=grp)
        packed = torch.stack([stats[:, 1] * torch.exp(local_max - gmax), stats[:, 2]])
        dist.all_reduce(packed, group=grp)
        lse = (gmax + packed[0].log()).contiguous()
        valid = targets != ignore_index
        cnt = valid.sum().clamp_min(1).float()
        loss = (lse - packed[1]).masked_fill(~valid, 0.0).sum() / cnt * grad_scale
        inv = (1.0 / cnt).reshape(1)
        if logits.requires_grad:
            native.call("llmctl_cross_entropy_finalize", logits.data_ptr(), logits.stride(0), targets.data_ptr(),
                        lse.data_ptr(), rows, v_local, int(v_start), int(ignore_index), float(grad_scale),
                        inv.data_ptr(), native.stream_ptr())
        ctx.save_for_backward(logits.detach())
        return loss

    @staticmethod
    def backward(ctx, dloss):
        (g,) = ctx.saved_tensors      # the logits buffer now holds d(loss)/d(logits); the loss is differentiated once
        return g, None, None, None, None


def vocab_parallel_cross_entropy(logits: torch.Tensor, targets: torch.Tensor, ignore_index: int = -100,
                                 grad_scale: float = 1.0) -> torch.Tensor:
    """Mean cross-entropy over logits sharded on the vocabulary.  On CUDA (native kernels) the logits buffer is
    consumed — overwritten with its gradient — and ``.backward()`` must be called on the returned loss itself."""
    st = get_state()
    V_local = logits.shape[-1]
    logits = logits.reshape(-1, V_local)
    targets = targets.reshape(-1)
    if st.tp == 1:
        return ops.cross_entropy(logits, targets, ignore_index, grad_scale=grad_scale, assume_unit_grad=True)
    if (ops._use_native(logits, "elementwise") and V_local % 8 == 0 and logits.stride(1) == 1
            and logits.stride(0) % 8 == 0 and not logits.is_leaf):
        return _VocabParallelCENative.apply(logits, targets.contiguous(), st.tp_rank * V_local, ignore_index, grad_scale)
    return _VocabParallelCE.apply(logits, targets, st.tp_rank * V_local, ignore_index, grad_scale)
