"""Rectified-flow schedule with the reference's API (model/rectified_flow.py).  The schedule itself
is scalar bookkeeping on tiny tensors; noise draws stay torch calls so seeds match the reference."""
import numpy as np
import torch


class RectifiedFlow:
    def __init__(self, num_steps=1000, theta=1e-5):
        self.num_step = num_steps
        self.theta = theta

    def get_train_tuple(self, z0=None, z1=None, time_step=None):
        """rectified_flow.py:18-40: z_t = t z1 + (1-t) z0, target = z1 - z0."""
        if time_step is None:
            t = torch.rand((z1.shape[0], 1))
        else:
            t = self.timestep_to_time(time_step)
        if z0 is None:
            z0 = torch.randn_like(z1)
        return t * z1 + (1.0 - t) * z0, z1 - z0

    def get_target_with_zt_vel(self, zt, vel, time_step):
        """rectified_flow.py:42-45."""
        return zt + (1.0 - self.timestep_to_time(time_step)) * vel

    @torch.no_grad()
    def timestep_to_time(self, time_step):
        """rectified_flow.py:70-75."""
        t = (self.num_step - time_step) / self.num_step
        if len(t.shape) == 1:
            t = t.view(-1, 1, 1, 1)
        return t

    @torch.no_grad()
    def sample_loop(self, model, video_start, sample_step, start_step=None, motion=None, motion_available_length=None,
                    reinforce_condition=False):
        """rectified_flow.py:78-116 (note: this loop includes step 0, unlike AMDModel_New.sample)."""
        if start_step is None:
            start_step = self.num_step
        step_seq = np.linspace(0, start_step, num=sample_step, endpoint=True, dtype=int)
        dt = 1.0 / sample_step
        sample = video_start
        for i in list(reversed(step_seq)):
            time_step = torch.ones((sample.shape[0],), device=sample.device) * int(i)
            zt = torch.cat((video_start, sample), dim=1) if reinforce_condition else sample
            pre = model(zt, time_step) if motion is None else model(motion, zt, time_step, motion_available_length)
            sample = sample + pre * dt
        return sample

    @torch.no_grad()
    def sample_step(self, model, zt, sample_step, start_step=None):
        """rectified_flow.py:119-138."""
        if start_step is None:
            start_step = self.num_step
        time_step = torch.ones((zt.shape[0], 1), device=zt.device) * start_step
        return zt + model(zt, time_step) * (1.0 / sample_step)
