"""Loading the reference's Lightning checkpoints.

``model.ckpt`` written by the reference's Trainer is a ``torch.save``d dict whose ``'state_dict'`` holds the LightningModule's
tensors: ``model.*`` is the IVP (``model.dyn_fun.*``, ``model.init_coordinates.param_map.*`` ...), other prefixes belong to
samplers / candidates.  ``robustness/eval_utils.py:92-107`` (hydra_conf_load_from_checkpoint_nonstrict) loads it with
``strict=False`` because the lazily created conv ``alpha`` (layers.py:80-81) and the eager one of layers_test.py:65 differ
between training and test classes.  The mirrored module tree has the same names, so loading is a key filter."""
from __future__ import annotations

import torch


def reference_state_dict(ckpt):
    """Accepts a path, a Lightning checkpoint dict ({'state_dict': ...}) or a bare state_dict; returns the bare state_dict."""
    if isinstance(ckpt, (str, bytes)) or hasattr(ckpt, '__fspath__'):
        ckpt = torch.load(ckpt, map_location='cpu', weights_only=False)
    if isinstance(ckpt, dict) and 'state_dict' in ckpt and isinstance(ckpt['state_dict'], dict):
        ckpt = ckpt['state_dict']
    return ckpt


def load_reference_checkpoint(module, ckpt, strict=False):
    """Load a reference checkpoint into ``module``: a ``fiode_b200.pl_modules.LyapunovLearning`` (keys used as they are) or
    a ``fiode_b200.models.IVP`` (the ``model.`` prefix is stripped).  Returns (missing, unexpected) like load_state_dict;
    with ``strict=False`` (the reference's own choice) only shape mismatches raise.  Cached Cayley matrices are dropped."""
    sd = reference_state_dict(ckpt)
    own = set(module.state_dict().keys()) | {n for n, _ in module.named_parameters()}
    if not any(k in own for k in sd) and any(k.startswith('model.') for k in sd):
        sd = {k[len('model.'):]: v for k, v in sd.items() if k.startswith('model.')}
    res = module.load_state_dict(sd, strict=strict)
    for m in module.modules():
        if hasattr(m, '_q_key'):
            m._q_key = None                          # weights changed: recompute Q / Qf on next use
    return res
