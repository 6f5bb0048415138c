"""`gsr_activations.activate`: the three parameter activations of scene/gaussian_model.py:43-51 / 95-106 --
get_opacity = sigmoid(_opacity), get_scaling = exp(_scaling), get_rotation = normalize(_rotation) -- as ONE kernel
per direction (csrc/activations.cuh) instead of ~8 torch kernels forward and ~17 in the
autograd replay.  Values are bit-identical to torch's on the same GPU, so the rasterizer sees the same inputs.

    opacity, scales, rotations = activate(pc._opacity, pc._scaling, pc._rotation)
"""
import torch

from diff_gaussian_rasterization import _lib


def _dense(t, shape):
    if not t.is_cuda:
        raise RuntimeError("gsr_activations: tensors must be CUDA tensors (there is no CPU path)")
    if tuple(t.shape) != shape:
        raise RuntimeError("gsr_activations: expected shape %s, got %s" % (shape, tuple(t.shape)))
    return t.detach().to(torch.float32).contiguous()


class _Activate(torch.autograd.Function):
    @staticmethod
    def forward(ctx, opacity_raw, scaling_raw, rotation_raw):
        P = opacity_raw.size(0)
        o, s, q = _dense(opacity_raw, (P, 1)), _dense(scaling_raw, (P, 3)), _dense(rotation_raw, (P, 4))
        out = torch.empty_like(o), torch.empty_like(s), torch.empty_like(q)
        if P:
            with torch.cuda.device(o.device):
                _lib.check(_lib.lib.gsr_activate(P, o.data_ptr(), s.data_ptr(), q.data_ptr(), out[0].data_ptr(),
                                                 out[1].data_ptr(), out[2].data_ptr(),
                                                 torch.cuda.current_stream(o.device).cuda_stream))
        ctx.save_for_backward(o, s, q)
        return out

    @staticmethod
    def backward(ctx, g_o, g_s, g_q):
        o, s, q = ctx.saved_tensors
        P = o.size(0)
        z = lambda g, like: torch.zeros_like(like) if g is None else g.to(torch.float32).contiguous()
        g_o, g_s, g_q = z(g_o, o), z(g_s, s), z(g_q, q)
        d = torch.empty_like(o), torch.empty_like(s), torch.empty_like(q)
        if P:
            with torch.cuda.device(o.device):
                _lib.check(_lib.lib.gsr_activate_backward(P, o.data_ptr(), s.data_ptr(), q.data_ptr(), g_o.data_ptr(),
                                                          g_s.data_ptr(), g_q.data_ptr(), d[0].data_ptr(),
                                                          d[1].data_ptr(), d[2].data_ptr(),
                                                          torch.cuda.current_stream(o.device).cuda_stream))
        return d


def activate(opacity_raw, scaling_raw, rotation_raw):
    """(sigmoid(opacity_raw) (P,1), exp(scaling_raw) (P,3), normalize(rotation_raw) (P,4)), differentiable."""
    return _Activate.apply(opacity_raw, scaling_raw, rotation_raw)
