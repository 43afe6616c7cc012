"""One decoder training step on the fused path (BASELINE configs[3]).

The shape of the reference's loop body (main/train_pano2gaussian_decoder.py:219-265): decode,
copy the five results into a GaussianModel, render, loss, backward, optimiser step — with the
gradient all-reduce of ``data_parallel.py`` between backward and step when several GPUs train
data-parallel.  The reference's extra losses (LPIPS / ID / Sobel) need pretrained networks and
are out of scope; L1 + L2 on the image are used (gaussian_splatting/utils/loss_utils.py:17-24).
"""
import torch

from .data_parallel import allreduce_gradients
from .gaussian_renderer import GaussianModel, render_simple


def l1_loss(a, b):
    return torch.abs(a - b).mean()


def l2_loss(a, b):
    return ((a - b) ** 2).mean()


def training_step(decoder, optimizer, z, gan_camera_params, init_position, camera, target, bg,
                  truncation_psi=1.0, group=None, dL_dimage=None):
    """Returns the loss tensor (device).  ``dL_dimage``: if given, backpropagate that image
    gradient instead of the L1+L2 loss (benchmark mode of SURVEY §8d, config 4)."""
    result = decoder(z, gan_camera_params, init_position, truncation_psi)
    gaussians = GaussianModel.from_decoder(result)        # activations through autograd
    image = render_simple(camera, gaussians, bg)["render"]
    if dL_dimage is not None:
        loss = (image * dL_dimage).sum()
    else:
        loss = l1_loss(image, target) + l2_loss(image, target)
    loss.backward()
    extra = list(decoder.G.parameters()) if decoder.use_gen_finetune else ()
    allreduce_gradients(decoder, extra_params=extra, group=group)
    optimizer.step()
    # Dropping the .grad tensors costs nothing; the next backward hands out fresh views of ONE flat gradient
    # buffer (fused_backward.py), so neither 40 fill kernels nor 40 accumulate kernels run per step.
    optimizer.zero_grad(set_to_none=True)
    return loss.detach()
