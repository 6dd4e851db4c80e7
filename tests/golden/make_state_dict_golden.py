"""Generates tests/golden/state_dict_reference.json by IMPORTING THE UNMODIFIED REFERENCE: the keys, shapes and dtypes
of GaussianDiffusion(TemporalUnet(...)).state_dict() -- what Trainer.save stores under 'model' / 'ema'
(diffuser/utils/training.py:133-137) -- plus the loss_fn.weights values.  Build container only:
    PYTHONDONTWRITEBYTECODE=1 python tests/golden/make_state_dict_golden.py
"""
import contextlib
import io
import json
import os
import sys

import torch

HERE = os.path.dirname(os.path.abspath(__file__))
sys.dont_write_bytecode = True
sys.path.insert(0, "/root/reference/diffusion")

with contextlib.redirect_stdout(io.StringIO()):
    from diffuser.models.diffusion import GaussianDiffusion  # noqa: E402
    from diffuser.models.temporal import TemporalUnet  # noqa: E402

out = {}
for name, kw, dkw in (
        ("default_h64", dict(horizon=64, transition_dim=35, cond_dim=35), dict(n_timesteps=100, predict_epsilon=False)),
        ("attention_h32_l2", dict(horizon=32, transition_dim=35, cond_dim=35, attention=True),
         dict(n_timesteps=50, loss_type="l2", action_weight=2.0, loss_discount=0.99)),
        ("wide_t69_dim128", dict(horizon=32, transition_dim=69, cond_dim=0, dim=128), dict(n_timesteps=20))):
    with contextlib.redirect_stdout(io.StringIO()):
        net = TemporalUnet(**kw)
        diff = GaussianDiffusion(net, kw["horizon"], kw["transition_dim"], **dkw)
    sd = diff.state_dict()
    out[name] = {"unet": kw, "diffusion": dkw,
                 "keys": {k: [list(v.shape), str(v.dtype)] for k, v in sd.items()},
                 "loss_fn.weights": sd["loss_fn.weights"].tolist()}
json.dump(out, open(os.path.join(HERE, "state_dict_reference.json"), "w"))
print({k: len(v["keys"]) for k, v in out.items()})
