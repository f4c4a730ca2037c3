"""Latent-space equivalent of the reference's inference pipeline (model/pipeline.py:620-751, AMDPipeLine_single): the
reference reads a video, VAE-encodes window + 1 frames, reconstructs the window from (reference frame, motion tokens) with
`amd_model.sample`, VAE-decodes and writes a video.  Video I/O and the VAE are outside the hot path: this class
takes and returns VAE latents; `vae_encode` / `vae_decode` callables may be plugged in for pixel-space use."""
import torch


class AMDPipeLine_single:
    def __init__(self, vae=None, amd_model=None, window=16, device="cuda:0", use_grey=False, vae_encode=None, vae_decode=None,
                 **kwargs):
        if use_grey:
            raise NotImplementedError("use_grey needs the grey-scale VAE latents of the reference pipeline (outside the path)")
        self.vae, self.amd_model, self.window, self.ddevice = vae, amd_model, window, device
        self.vae_encode, self.vae_decode = vae_encode, vae_decode

    @torch.no_grad()
    def sample_latents(self, videos_z, video_sample_step=20, drop_prev_img=False):
        """videos_z (window + 1, C, H, W): frame 0 is the reference image (pipeline.py:684-725).
        Returns (1, window + 1, C, H, W) = [reference | reconstruction]."""
        W = self.window
        assert videos_z.shape[0] == W + 1, f"need window + 1 = {W + 1} latent frames, got {videos_z.shape[0]}"
        videos_z = videos_z.to(self.ddevice)
        refimg_z = videos_z[:1]
        gt_video_z = videos_z[1:].unsqueeze(0)
        prev = refimg_z.unsqueeze(1).repeat(1, W, 1, 1, 1)
        if drop_prev_img:
            prev = torch.zeros_like(prev)
        _, video_pre, _ = self.amd_model.sample(video=gt_video_z.contiguous(), ref_img=prev.contiguous(), sample_step=video_sample_step)
        return torch.cat([refimg_z.unsqueeze(1), video_pre], dim=1)

    @torch.no_grad()
    def sample_long(self, videos_z, video_sample_step=20, chain=False):
        """A latent sequence longer than one window, cut into consecutive windows that are reconstructed in ONE batched
        `sample` call (the clips of a batch are independent).  Window k covers frames [k W + 1, (k + 1) W + 1) and is
        conditioned on frame k W of the input (chain=False) — or, chain=True, on the last frame reconstructed for window
        k - 1 (sequential, the `cur_prev_img = ...[S-1]` alternative commented in the reference, pipeline.py:707)."""
        W = self.window
        F = videos_z.shape[0]
        assert (F - 1) % W == 0 and F > W, f"need k * window + 1 latent frames, got {F}"
        videos_z = videos_z.to(self.ddevice)
        K = (F - 1) // W
        gt = videos_z[1:].view(K, W, *videos_z.shape[1:])
        if not chain:
            ref = videos_z[0:F - 1:W].unsqueeze(1).repeat(1, W, 1, 1, 1)
            _, pre, _ = self.amd_model.sample(video=gt.contiguous(), ref_img=ref.contiguous(), sample_step=video_sample_step)
            return torch.cat([videos_z[:1], pre.flatten(0, 1)], 0).unsqueeze(0)
        out, prev = [videos_z[:1]], videos_z[:1]
        for k in range(K):
            ref = prev.unsqueeze(1).repeat(1, W, 1, 1, 1)
            _, pre, _ = self.amd_model.sample(video=gt[k:k + 1].contiguous(), ref_img=ref.contiguous(), sample_step=video_sample_step)
            out.append(pre[0])
            prev = pre[0, -1:]
        return torch.cat(out, 0).unsqueeze(0)

    @torch.no_grad()
    def sample(self, videos, video_sample_step=20, **kwargs):
        """pixel-space entry (needs the vae_encode / vae_decode callables; the VAE itself is outside this package)"""
        if self.vae_encode is None or self.vae_decode is None:
            raise RuntimeError("AMDPipeLine_single.sample needs vae_encode / vae_decode callables; use sample_latents for VAE latents")
        z = self.vae_encode(self.vae, videos)
        return self.vae_decode(self.vae, self.sample_latents(z, video_sample_step)[0])
