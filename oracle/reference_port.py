"""ORACLE -- TEST INFRASTRUCTURE ONLY.  Never imported by the product package.

CPU restatement (plain PyTorch, dense, fp32 or fp64) of the reference's GraphCNNDropEdge hot path, written
as state-dict-driven functions with injectable dropout masks.  Only ``tests/``, ``__graft_entry__.smoke()``
and ``bench.py``'s ``cpu_baseline`` / ``--impl reference`` legs may import this module.

Parity status: the reference's own tests hold no vectors for ``gnn/``, so this port is
pinned against OUTPUTS OF THE REFERENCE ITSELF: ``oracle/make_golden.py`` imports the unmodified reference from
``/root/reference`` (authoring container only), runs it on seeded inputs with injected masks and commits the
vectors under ``tests/golden/``; ``tests/test_oracle.py`` checks this port against those vectors (and, when
``/root/reference`` is present, against the live reference).

Every function cites the reference lines it restates (paths relative to the reference root).
"""
from __future__ import annotations

from typing import Dict, List, Optional, Sequence, Tuple

import torch

P_FEAT = 0.5   # nn.Dropout(p=0.5)  gnn/models/networks/drop_robust_gcn.py:37
P_EDGE = 0.3   # nn.Dropout(p=0.3)  gnn/models/networks/drop_robust_gcn.py:38


def preprocess_adj(adj_bijl: torch.Tensor) -> torch.Tensor:
    """gnn/models/networks/robust_gcn.py:53-72.  adj_bijl: B x N x N x L  ->  A': B x (L+1)N x N with
    A'[b, i*(L+1)+t, j]; t = 0 is the identity block."""
    B, N, _, L = adj_bijl.shape
    eye = torch.eye(N, dtype=adj_bijl.dtype).reshape(1, N, N, 1).expand(B, N, N, 1)
    stacked = torch.cat([eye, adj_bijl], dim=-1)                      # B x N x N x (L+1)      (:65)
    stacked = stacked.reshape(B * N, N, L + 1)                        # (:66)
    return stacked.transpose(1, 2).reshape(B, (L + 1) * N, N)         # (:71)


def graph_conv(V: torch.Tensor, A_prime: torch.Tensor, h_weights: torch.Tensor, bias: torch.Tensor) -> torch.Tensor:
    """gnn/models/networks/robust_gcn.py:45-51: aggregate (A' V) then project with K = (L+1)F."""
    B, N, F = V.shape
    K = h_weights.shape[0]
    new_v = torch.matmul(A_prime, V).reshape(B, N, K)                 # (:45-47)
    return torch.matmul(new_v, h_weights) + bias.unsqueeze(0)         # (:50)


def node_self_atten(V: torch.Tensor, sd: Dict[str, torch.Tensor], prefix: str = "self_atten.") -> torch.Tensor:
    """gnn/models/networks/robust_gcn.py:90-96: unscaled, unmasked dense attention; gamma * o + V."""
    def lin_relu(name):
        return torch.relu(torch.nn.functional.linear(V, sd[prefix + name + ".0.weight"], sd[prefix + name + ".0.bias"]))
    f_out, g_out, h_out = lin_relu("f"), lin_relu("g"), lin_relu("h")
    s = torch.softmax(torch.matmul(f_out, g_out.transpose(1, 2)), dim=-1)
    return sd[prefix + "gamma"] * torch.matmul(s, h_out) + V


def _drop(x: torch.Tensor, p: float, training: bool, mask: Optional[torch.Tensor]) -> torch.Tensor:
    """Inverted dropout as nn.Dropout does it; ``mask`` (keep bits) replaces the Bernoulli draw."""
    if not training or p == 0.0:
        return x
    if mask is None:
        return torch.nn.functional.dropout(x, p, True)
    return x * mask.to(x.dtype) / (1.0 - p)


def forward(sd: Dict[str, torch.Tensor], V: torch.Tensor, A: torch.Tensor, training: bool = False,
            masks: Optional[Dict[str, Sequence[torch.Tensor]]] = None, use_attention: bool = True) -> torch.Tensor:
    """gnn/models/networks/drop_robust_gcn.py:61-103.

    V: B x N x input_dim; A: B x N x L x N ([b, start, type, end]) -- both already ``.float()``-ed as
    gnn/trainer/training_procedures/kv_procedure.py:116-119 does.
    masks: {"edge": 3 keep-masks B x (L+1)N x N, "feat": 5 keep-masks (4 x B x N x S, 1 x B x N x 10H)}.
    """
    em = masks["edge"] if masks else [None] * 3
    fm = masks["feat"] if masks else [None] * 5
    lin = torch.nn.functional.linear
    A = A.permute(0, 1, 3, 2)                                                           # (:63)
    emb = _drop(torch.relu(lin(V, sd["emb1.0.weight"], sd["emb1.0.bias"])), P_FEAT, training, fm[0])   # (:64)
    Ap = preprocess_adj(A)                                                              # (:69)
    g1 = torch.relu(graph_conv(emb, _drop(Ap, P_EDGE, training, em[0]), sd["gcn1.h_weights"], sd["gcn1.bias"]))
    g1 = _drop(g1, P_FEAT, training, fm[1])                                             # (:76-77)
    g2 = torch.relu(graph_conv(g1, _drop(Ap, P_EDGE, training, em[1]), sd["gcn2.h_weights"], sd["gcn2.bias"]))
    g2 = _drop(g2, P_FEAT, training, fm[2])                                             # (:80-81)
    g3 = torch.relu(graph_conv(torch.cat([g1, g2], dim=-1), _drop(Ap, P_EDGE, training, em[2]),
                               sd["gcn3.h_weights"], sd["gcn3.bias"]))
    g3 = _drop(g3, P_FEAT, training, fm[3])                                             # (:84-86)
    new_v = torch.relu(lin(torch.cat([g1, g3], dim=-1), sd["emb2.0.weight"], sd["emb2.0.bias"]))     # (:88-89)
    if use_attention:
        new_v = node_self_atten(new_v, sd)                                              # (:94-95)
    new_v = _drop(torch.relu(lin(new_v, sd["w_rand.projection.weight"])), P_FEAT, training, fm[4])   # (:99-100)
    return lin(new_v, sd["classifier.weight"], sd["classifier.bias"])                   # (:103)


def cross_entropy(pred: torch.Tensor, target: torch.Tensor) -> torch.Tensor:
    """gnn/trainer/losses/cross_entropy_loss.py:33-35: CE over pred.transpose(1, 2); ignore_index = -100."""
    return torch.nn.functional.cross_entropy(pred.transpose(1, 2), target)


TRAINABLE_EXCLUDE = ("w_rand.projection.weight",)   # frozen: drop_robust_gcn.py:20-21


def loss_and_grads(sd: Dict[str, torch.Tensor], V: torch.Tensor, A: torch.Tensor, y: torch.Tensor,
                   training: bool = True, masks=None, use_attention: bool = True
                   ) -> Tuple[torch.Tensor, torch.Tensor, Dict[str, torch.Tensor]]:
    """forward + CE + backward (kv_procedure.py:116-128,159).  Returns (logits, loss, grads by state-dict key)."""
    leaves = {k: v.detach().clone().requires_grad_(k not in TRAINABLE_EXCLUDE) for k, v in sd.items()}
    logits = forward(leaves, V, A, training, masks, use_attention)
    loss = cross_entropy(logits, y)
    names = [k for k in leaves if leaves[k].requires_grad]
    grads = torch.autograd.grad(loss, [leaves[k] for k in names])
    return logits.detach(), loss.detach(), dict(zip(names, grads))


def clip_grad_norm_(grads: Dict[str, torch.Tensor], max_norm: float) -> torch.Tensor:
    """torch.nn.utils.clip_grad_norm_ as called at kv_procedure.py:161 (L2, eps 1e-6, coefficient clamped to 1)."""
    total = torch.linalg.vector_norm(torch.stack([torch.linalg.vector_norm(g) for g in grads.values()]))
    coef = torch.clamp(max_norm / (total + 1e-6), max=1.0)
    for g in grads.values():
        g.mul_(coef)
    return total


def adam_step(sd: Dict[str, torch.Tensor], grads: Dict[str, torch.Tensor], state: Dict[str, Dict[str, torch.Tensor]],
              step: int, lr: float = 1e-3, betas=(0.9, 0.999), eps: float = 1e-8) -> None:
    """torch.optim.Adam defaults (the reference swallows the configured lr: builtin_optimizer.py:10,18-23)."""
    for k, g in grads.items():
        st = state.setdefault(k, {"m": torch.zeros_like(g), "v": torch.zeros_like(g)})
        st["m"].mul_(betas[0]).add_(g, alpha=1 - betas[0])
        st["v"].mul_(betas[1]).addcmul_(g, g, value=1 - betas[1])
        bc1, bc2 = 1 - betas[0] ** step, 1 - betas[1] ** step
        denom = (st["v"].sqrt() / (bc2 ** 0.5)).add_(eps)
        sd[k] = sd[k] - (lr / bc1) * st["m"] / denom


def train_step(sd, V, A, y, opt_state, step: int, max_grad_norm: float = 5.0, masks=None, use_attention=True):
    """KVProcedure._run_train_step (kv_procedure.py:143-164) without the sklearn report:
    zero_grad -> forward -> CE -> backward -> clip_grad_norm_ -> Adam.step.  Mutates ``sd`` in place."""
    logits, loss, grads = loss_and_grads(sd, V, A, y, True, masks, use_attention)
    clip_grad_norm_(grads, max_grad_norm)
    adam_step(sd, grads, opt_state, step)
    return logits, loss, grads
