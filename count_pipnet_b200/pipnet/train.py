"""Training step of PIP-Net / Count-PIP-Net behind the reference's API (pipnet/train.py of TarasKutsyk/Count_PIPNet):
`train_pipnet(...)`, `calculate_loss(...)`, `align_loss`, `uniform_loss` with the reference's argument lists and
return values.

What runs on the B200 here (first slice): for CUDA tensors `calculate_loss` is ONE autograd
node over `cpn_pipnet_loss` -- align, tanh and classification terms, their weighted sum, the accuracy count and every
gradient of the loss (w.r.t. the prototype map, the pooled scores, the logits and the classifier's
normalization_multiplier) from a single pass over the [2B, P, H, W] map; `softmax_pool_head` is the autograd node for
Softmax(dim=1) + spatial max / sum pooling on that map (`cpn_softmax_pool_nchw` / `_bwd`). The backbone under autograd is
still the composed PyTorch modules. For CPU tensors `calculate_loss` composes the same arithmetic from torch ops, like
the modules' autograd branch does; CUDA tensors never take that branch (a missing library raises).
"""
from __future__ import annotations

import torch
import torch.nn.functional as F

from .. import ops
from .._lib import POOL_MAX, POOL_SUM


# ---- autograd nodes over the C ABI ------------------------------------------------------------------------------------
class _FusedPipnetLoss(torch.autograd.Function):
    """loss, result[8] = cpn_pipnet_loss(...); the gradients are computed in the same launch and kept for backward."""

    @staticmethod
    def forward(ctx, proto_features, pooled, out, norm_mult, ys, cfg):
        align_w, tanh_w, class_w, tanh_coeff, eps, pretrain, finetune, use_log1p = cfg
        pf = proto_features.detach().contiguous()
        pl = pooled.detach().contiguous()
        if pf.dtype != torch.float32 or pl.dtype != torch.float32:
            raise TypeError("calculate_loss: the B200 kernels take fp32 proto_features / pooled")
        n2, P = pf.shape[0], pf.shape[1]
        if n2 % 2 or pl.shape != (n2, P):
            raise ValueError(f"calculate_loss: proto_features {tuple(pf.shape)} / pooled {tuple(pl.shape)} are not two "
                             "stacked views of one batch")
        B, HW = n2 // 2, pf[0, 0].numel()
        dev = pf.device
        need = ctx.needs_input_grad
        o = m = y = None
        K = 0
        if not pretrain:
            o = out.detach().contiguous().float()
            K = o.shape[1]
            m = norm_mult.detach().reshape(1).contiguous().float()
            y = ys.to(device=dev, dtype=torch.int64).contiguous()
            if y.numel() != B or o.shape[0] != n2:
                raise ValueError("calculate_loss: out / ys do not match the batch")
        g_pf = torch.empty_like(pf) if (need[0] and not finetune) else None
        g_pl = torch.empty_like(pl) if (need[1] and not finetune) else None
        g_o = torch.empty_like(o) if (o is not None and (need[2] or need[3])) else None
        result = torch.empty(8, device=dev, dtype=torch.float32)
        ws = torch.empty(ops.pipnet_loss_ws_floats(B, P, HW), device=dev, dtype=torch.float32)
        ops.pipnet_loss(pf, pl, o, y, m, B, P, HW, K, align_w, tanh_w, class_w, tanh_coeff, eps, pretrain, finetune,
                        use_log1p, g_pf, g_pl, g_o, result, ws)
        ctx.has = (g_pf is not None, g_pl is not None, g_o is not None)
        ctx.shapes = (proto_features.shape, pooled.shape, None if out is None else out.shape,
                      None if norm_mult is None else norm_mult.shape)
        ctx.save_for_backward(*[t for t in (g_pf, g_pl, g_o) if t is not None], result)
        ctx.mark_non_differentiable(result)
        return result[3].clone(), result

    @staticmethod
    def backward(ctx, g_loss, _g_result):
        saved = list(ctx.saved_tensors)
        result = saved.pop()
        g = g_loss.detach().reshape(1).float().contiguous()
        grads = []
        for present in ctx.has:
            t = saved.pop(0) if present else None
            if t is not None:
                ops.scale_unless_one(t, g)  # loss.backward() arrives with 1: one empty launch
            grads.append(t)
        g_pf, g_pl, g_o = grads
        need = ctx.needs_input_grad
        g_m = None
        if need[3] and g_o is not None:
            g_m = (result[5:6] * g).reshape(ctx.shapes[3])
        return (g_pf.view(ctx.shapes[0]) if (g_pf is not None and need[0]) else None,
                g_pl.view(ctx.shapes[1]) if (g_pl is not None and need[1]) else None,
                g_o.view(ctx.shapes[2]) if (g_o is not None and need[2]) else None, g_m, None, None)


class _SoftmaxPoolHead(torch.autograd.Function):
    """(proto_features, pooled) = softmax over dim 1 of an NCHW map + spatial max / sum; backward is one kernel."""

    @staticmethod
    def forward(ctx, z, mode):
        zc = z.detach().contiguous()
        if zc.dtype != torch.float32 or zc.dim() != 4:
            raise TypeError("softmax_pool_head: expects an fp32 [N, P, H, W] map")
        N, P = zc.shape[0], zc.shape[1]
        HW = zc.shape[2] * zc.shape[3]
        pf = torch.empty_like(zc)
        pooled = torch.empty(N, P, device=zc.device, dtype=torch.float32)
        argmax = torch.empty(N, P, device=zc.device, dtype=torch.int32) if mode == POOL_MAX else None
        ws = torch.empty(ops.softmax_pool_nchw_ws_floats(N, P, HW), device=zc.device, dtype=torch.float32)
        ops.softmax_pool_nchw(zc, mode, pf, pooled, argmax, ws)
        ctx.mode = mode
        ctx.set_materialize_grads(False)  # an unused output arrives as None, not as a map of zeros to be read
        if argmax is not None:
            ctx.save_for_backward(pf, argmax)
        else:
            ctx.save_for_backward(pf)
        return pf, pooled

    @staticmethod
    def backward(ctx, g_pf, g_pooled):
        pf = ctx.saved_tensors[0]
        argmax = ctx.saved_tensors[1] if ctx.mode == POOL_MAX else None
        g_pf = None if g_pf is None else g_pf.contiguous()
        g_pooled = None if g_pooled is None else g_pooled.contiguous()
        grad_z = torch.empty_like(pf)
        ops.softmax_pool_nchw_bwd(pf, g_pf, g_pooled, argmax, ctx.mode, grad_z)
        return grad_z, None


def softmax_pool_head(z: torch.Tensor, mode: str = "max"):
    """Softmax(dim=1) then AdaptiveMaxPool2d((1, 1)) + Flatten (pipnet/pipnet.py:96, 105-108; mode "max") or the spatial
    sum (pipnet/count_pipnet.py:88; mode "sum") of a CUDA [N, P, H, W] map, differentiable: -> (proto_features, pooled)."""
    if not z.is_cuda:
        raise ops._lib.CpnError("softmax_pool_head runs on a B200 only; compose nn.Softmax and the pooling on CPU tensors")
    return _SoftmaxPoolHead.apply(z, POOL_MAX if mode == "max" else POOL_SUM)


# ---- pipnet/train.py ---------------------------------------------------------------------------------------------------
def align_loss(inputs, targets, EPS=1e-12):
    """-mean log(<inputs_n, targets_n> + EPS) over the rows (pipnet/train.py:253-265)."""
    assert inputs.shape == targets.shape
    assert targets.requires_grad == False  # noqa: E712  (the reference's own check)
    dots = (inputs * targets).sum(dim=1)
    return -torch.log(dots + EPS).mean()


def uniform_loss(x, t=2):
    """log mean exp(-t * squared pairwise distance) (pipnet/train.py:247-250; unused by the reference's step too)."""
    return (torch.pdist(x, p=2).pow(2).mul(-t).exp().mean() + 1e-10).log()


def _composed_loss(proto_features, pooled, out, ys1, align_pf_weight, t_weight, cl_weight, norm_mult, pretrain, finetune,
                   criterion, coeff, EPS, enforce_weight_sparsity):
    """The same terms from torch ops, for tensors that are not on a GPU (pipnet/train.py:165-229)."""
    ys = torch.cat([ys1, ys1])
    pooled1, pooled2 = pooled.chunk(2)
    pf1, pf2 = proto_features.chunk(2)
    e1 = pf1.flatten(start_dim=2).permute(0, 2, 1).flatten(end_dim=1)
    e2 = pf2.flatten(start_dim=2).permute(0, 2, 1).flatten(end_dim=1)
    a_loss = (align_loss(e1, e2.detach()) + align_loss(e2, e1.detach())) / 2.
    tanh_loss = -(torch.log(torch.tanh(torch.sum(coeff * pooled1, dim=0)) + EPS).mean()
                  + torch.log(torch.tanh(torch.sum(coeff * pooled2, dim=0)) + EPS).mean()) / 2.
    loss = None
    class_loss = None
    if not finetune:
        loss = align_pf_weight * a_loss + t_weight * tanh_loss
    correct = 0
    if not pretrain:
        inputs = torch.log1p(out ** norm_mult) if enforce_weight_sparsity else out
        class_loss = criterion(F.log_softmax(inputs, dim=1), ys)
        loss = cl_weight * class_loss if finetune else loss + cl_weight * class_loss
        correct = torch.sum(torch.eq(torch.argmax(out, dim=1), ys)).item()
    return loss, a_loss.item(), tanh_loss.item(), (0.0 if class_loss is None else class_loss.item()), correct


def calculate_loss(proto_features, pooled, out, ys1, align_pf_weight, t_weight, unif_weight, cl_weight,
                   net_normalization_multiplier, pretrain, finetune, criterion, train_iter,
                   is_count_pipnet=False, verbose=True, EPS=1e-10, enforce_weight_sparsity=True,
                   net=None, tanh_loss_coeff=1.0):
    """pipnet/train.py:165-243: returns (loss, acc, loss_components). `criterion` must be the reference's
    nn.NLLLoss(reduction='mean') (main.py:208) on the GPU path, which computes it inside the kernel."""
    coeff = float(tanh_loss_coeff) if is_count_pipnet else 1.0
    if is_count_pipnet and verbose:
        print(f'tanh_loss_coeff = {coeff}', flush=True)
        print('Using standard linear normalization')
    if proto_features.is_cuda:
        if not pretrain and criterion is not None:
            if not isinstance(criterion, torch.nn.NLLLoss) or criterion.reduction != 'mean' \
                    or criterion.weight is not None or criterion.ignore_index >= 0:
                raise ValueError("calculate_loss (B200 path): criterion must be nn.NLLLoss(reduction='mean') without class "
                                 "weights, as in the reference's main.py:208")
        cfg = (float(align_pf_weight), float(t_weight), float(cl_weight), coeff, float(EPS), bool(pretrain),
               bool(finetune), bool(enforce_weight_sparsity))
        loss, result = _FusedPipnetLoss.apply(proto_features, pooled, None if pretrain else out,
                                              None if pretrain else net_normalization_multiplier,
                                              None if pretrain else ys1, cfg)
        a_val, t_val, c_val, _, correct, _, _, _ = result.tolist()  # the step's one device-to-host read
    else:
        loss, a_val, t_val, c_val, correct = _composed_loss(
            proto_features, pooled, out, ys1, align_pf_weight, t_weight, cl_weight, net_normalization_multiplier,
            pretrain, finetune, criterion, coeff, EPS, enforce_weight_sparsity)
    loss_components = {
        'align': a_val,
        'align_weighted': a_val * align_pf_weight,
        'tanh': t_val,
        'tanh_weighted': t_val * t_weight,
        'class': 0.0 if pretrain else c_val,
        'class_weighted': 0.0 if pretrain else c_val * cl_weight,
    }
    acc = 0.
    if not pretrain:
        acc = correct / float(2 * len(ys1))
    if verbose and train_iter is not None and hasattr(train_iter, 'set_postfix_str'):
        with torch.no_grad():
            nz = torch.count_nonzero(torch.relu(pooled - 0.1), dim=1).float().mean().item()
            if pretrain:
                msg = f'L: {loss.item():.3f}, LA:{a_val:.2f}, LT:{t_val:.3f}, num_scores>0.1:{nz:.1f}'
            else:
                msg = (f'L:{loss.item():.3f},LC:{c_val:.3f}, LA:{a_val:.2f}, LT:{t_val:.3f}, '
                       f'num_scores>0.1:{nz:.1f}, Ac:{acc:.3f}')
            train_iter.set_postfix_str(msg, refresh=False)
    return loss, acc, loss_components


def _unwrapped(net):
    return net.module if hasattr(net, 'module') else net


def train_pipnet(net, train_loader, optimizer_net, optimizer_classifier, scheduler_net, scheduler_classifier, criterion,
                 epoch, nr_epochs, device, is_count_pipnet=False, pretrain=False, finetune=False,
                 progress_prefix: str = 'Train Epoch', enforce_weight_sparsity=True, tanh_loss_coeff=1.0):
    """One epoch (pipnet/train.py:8-163). `net` may be the bare model or wrapped (`net.module`, reference main.py:118)."""
    from tqdm import tqdm

    net.train()
    model = _unwrapped(net)
    cls = model._classification
    if pretrain:
        cls.requires_grad = False  # the reference sets this module attribute too (train.py:18); the optimizers decide
        progress_prefix = 'Pretrain Epoch'
    else:
        cls.requires_grad = True

    totals = dict.fromkeys(('align', 'tanh', 'class', 'align_weighted', 'tanh_weighted', 'class_weighted'), 0.)
    total_loss = 0.
    total_acc = 0.
    iters = len(train_loader)
    train_iter = tqdm(enumerate(train_loader), total=iters, desc=progress_prefix + '%s' % epoch, mininterval=2., ncols=0)
    print("Number of parameters that require gradient: ", sum(1 for p in net.parameters() if p.requires_grad), flush=True)

    if pretrain:
        align_pf_weight, unif_weight, t_weight, cl_weight = (epoch / nr_epochs) * 1., 0.5, 5., 0.
    else:
        align_pf_weight, unif_weight, t_weight, cl_weight = 5., 0., 2., 2.
    print("Align weight: ", align_pf_weight, ", U_tanh weight: ", t_weight, "Class weight:", cl_weight, flush=True)
    print("Pretrain?", pretrain, "Finetune?", finetune, flush=True)

    lrs_net, lrs_class = [], []
    i = -1
    for i, (xs1, xs2, ys) in train_iter:
        xs1, xs2, ys = xs1.to(device), xs2.to(device), ys.to(device)
        optimizer_classifier.zero_grad(set_to_none=True)
        optimizer_net.zero_grad(set_to_none=True)
        proto_features, pooled, out = net(torch.cat([xs1, xs2]))
        loss, acc, comps = calculate_loss(
            proto_features, pooled, out, ys, align_pf_weight, t_weight, unif_weight, cl_weight,
            cls.normalization_multiplier, pretrain, finetune, criterion, train_iter, is_count_pipnet=is_count_pipnet,
            EPS=1e-8, enforce_weight_sparsity=enforce_weight_sparsity, net=net if is_count_pipnet else None,
            tanh_loss_coeff=tanh_loss_coeff, verbose=(i == 0))
        for k in totals:
            totals[k] += comps[k]
        loss.backward()
        if not pretrain:
            optimizer_classifier.step()
            scheduler_classifier.step(epoch - 1 + (i / iters))
            lrs_class.append(scheduler_classifier.get_last_lr()[0])
        if not finetune:
            optimizer_net.step()
            scheduler_net.step()
            lrs_net.append(scheduler_net.get_last_lr()[0])
        else:
            lrs_net.append(0.)
        with torch.no_grad():
            total_acc += acc
            total_loss += loss.item()
            if not pretrain and enforce_weight_sparsity:
                # classifier weights below 1e-3 go to zero, multiplier >= 1, bias >= 0 (train.py:132-138)
                cls.weight.copy_(torch.clamp(cls.weight.data - 1e-3, min=0.))
                cls.normalization_multiplier.copy_(torch.clamp(cls.normalization_multiplier.data, min=1.0))
                if cls.bias is not None:
                    cls.bias.copy_(torch.clamp(cls.bias.data, min=0.))

    n = float(i + 1)
    train_info = {
        'align_loss_raw': totals['align'] / n, 'tanh_loss_raw': totals['tanh'] / n, 'class_loss_raw': totals['class'] / n,
        'align_loss_weighted': totals['align_weighted'] / n, 'tanh_loss_weighted': totals['tanh_weighted'] / n,
        'class_loss_weighted': totals['class_weighted'] / n,
    }
    print(f"\nEpoch {epoch} loss breakdown:")
    print(f"  Alignment loss: {train_info['align_loss_raw']:.4f} (raw), {train_info['align_loss_weighted']:.4f} (weighted)")
    print(f"  Tanh loss: {train_info['tanh_loss_raw']:.4f} (raw), {train_info['tanh_loss_weighted']:.4f} (weighted)")
    print(f"  Classification loss: {train_info['class_loss_raw']:.4f} (raw), {train_info['class_loss_weighted']:.4f} (weighted)")
    train_info['train_accuracy'] = total_acc / n
    train_info['loss'] = total_loss / n
    train_info['lrs_net'] = lrs_net
    train_info['lrs_class'] = lrs_class
    return train_info
