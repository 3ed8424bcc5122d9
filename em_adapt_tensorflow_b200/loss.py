"""The EM-Adapt loss as one differentiable torch op (deeplab.py:109-124 + 180-186).

The reference wires `tf.py_func(estep) -> tf.argmax -> sparse_softmax_cross_entropy_with_logits`
over the fc8 tensor; neither py_func nor argmax carries a gradient, so the labels are constants
and d loss / d fc8 = (softmax(fc8) - onehot(labels)) / n_pixels.  Here the whole chain stays on the
device: label sets from the (full-resolution) ground truth, the E-step, the labels and the loss with
its gradient are queued on the current stream by libweaklabel_cuda.so; `torch.autograd` only sees one
Function whose backward hands out the gradient the forward already computed.  No CPU path.
"""
import torch

from .plan import EStepPlan, OrderStream


class _EStepCrossEntropy(torch.autograd.Function):
    @staticmethod
    def forward(ctx, fc8, plan):
        labels, loss, valid, grad = plan.labels_and_loss(fc8, want_grad=True)
        ctx.save_for_backward(grad)
        ctx.mark_non_differentiable(labels)
        return loss, labels

    @staticmethod
    def backward(ctx, g_loss, _g_labels):
        (grad,) = ctx.saved_tensors
        return grad * g_loss, None


class EMAdaptLoss(torch.nn.Module):
    """loss = mean softmax-CE(fc8, argmax E-step(fc8 | image-level label sets)) -- ADAPT.getloss's
    `self.loss["norm"]` (deeplab.py:180-183) with the reference's constants as defaults (:181).

    forward(fc8, gt): fc8 CUDA float32 [B,H,W,C] (NHWC, contiguous), gt the dense ground-truth map
    [B,Hg,Wg] (uint8 / int32, CUDA) at any resolution -- only its image-level label set is used
    (weak supervision).  Returns (loss, labels uint8 [B,H,W])."""

    def __init__(self, bg_p=0.4, fg_p=0.2, num_iter=5, suppress_others=True, margin_others=1e-5, seed=None):
        super().__init__()
        self.args = dict(suppress_others=suppress_others, num_iter=num_iter, margin_others=margin_others, bg=bg_p, fg=fg_p)
        self.rng = OrderStream(seed) if seed is not None else None     # None: the process-global rand(), as the reference
        self._plans = {}

    def plan_for(self, fc8):
        key = (tuple(fc8.shape), fc8.device.index)
        if key not in self._plans:
            B, H, W, C = fc8.shape
            self._plans[key] = EStepPlan(B, H, W, C, device=fc8.device.index)
        return self._plans[key]

    def forward(self, fc8, gt):
        assert fc8.is_cuda and fc8.dim() == 4, "fc8 must be a CUDA tensor [B,H,W,C] (there is no CPU path)"
        fc8 = fc8.contiguous()
        plan = self.plan_for(fc8)
        plan.prepare_from_gt(gt.contiguous(), rng=self.rng, **self.args)
        return _EStepCrossEntropy.apply(fc8, plan)


def em_adapt_loss(fc8, gt, module=None, **kw):
    """Functional form; keeps one module (plans, order stream) per argument set."""
    if module is None:
        key = tuple(sorted(kw.items()))
        module = em_adapt_loss._cache.setdefault(key, EMAdaptLoss(**kw))
    return module(fc8, gt)


em_adapt_loss._cache = {}
