"""EMA of the weights (oracle; test infrastructure only).

The reference keeps ``diffusers.training_utils.EMAModel(self.model.parameters(), decay=0.999)``
(common/trainer.py:264-268) and calls ``ema_model.step(self.model.parameters())`` after every optimiser step
(common/trainer.py:350-351).  diffusers is not vendored in /root/reference and not installed, so this
is its published algorithm restated with the SAME torch tensor ops, in the reference's order
(parity unpinned by the reference: it holds no fixture for it):

    EMAModel.__init__   shadow_params = [p.clone().detach() for p in parameters]; optimization_step = 0
    EMAModel.get_decay  step = max(0, optimization_step - update_after_step - 1); 0.0 if step <= 0
                        cur = (1 + step) / (10 + step)      (use_ema_warmup False, the reference's setting)
                        min(cur, decay) clamped below by min_decay
    EMAModel.step       optimization_step += 1; decay = get_decay(optimization_step); omd = 1 - decay
                        trainable: s_param.sub_(omd * (s_param - param));  frozen: s_param.copy_(param)
"""
import torch


def get_decay(optimization_step, decay=0.999, min_decay=0.0, update_after_step=0):
    step = max(0, optimization_step - update_after_step - 1)
    if step <= 0:
        return 0.0
    cur = (1 + step) / (10 + step)
    return max(min(cur, decay), min_decay)


class EMAOracle:
    def __init__(self, parameters, decay=0.999):
        self.shadow_params = [p.clone().detach() for p in parameters]
        self.decay = decay
        self.optimization_step = 0

    @torch.no_grad()
    def step(self, parameters):
        self.optimization_step += 1
        one_minus_decay = 1 - get_decay(self.optimization_step, self.decay)
        for s_param, param in zip(self.shadow_params, parameters):
            if param.requires_grad:
                s_param.sub_(one_minus_decay * (s_param - param))       # three torch ops, each rounded to the dtype
            else:
                s_param.copy_(param)
        return one_minus_decay
