"""Named constructor-kwarg sets for AMDModel_New (model_AMD.py:1667-1721)."""

# scripts/amd_new.sh:37-67 + AMD_N factory (model_AMD.py:3151-3161): 695 M parameters.
DEFAULT = dict(
    image_inchannel=4, image_height=32, image_width=32, video_frames=16, scheduler_num_step=1000,
    use_filter=True, use_grey=False, use_camera_down=True, use_regularizers=False, use_camera=True, use_object=True,
    object_motion_token_num=8, object_motion_token_channel=16, object_enc_num_layers=8,
    camera_motion_token_num=8, camera_motion_token_channel=8, camera_enc_num_layers=8,
    enc_nhead=8, enc_ndim=64, motion_need_norm_out=False, motion_token_num=40, need_motion_transformer=False,
    diffusion_model_type="spatial", diffusion_attn_head_dim=64, diffusion_attn_num_heads=16,
    diffusion_out_channels=4, diffusion_num_layers=12, image_patch_size=2, motion_patch_size=1,
)

# SURVEY §8d config 1 ("Small", 14.1 M parameters) — the CPU-runnable parity case.
SMALL = dict(DEFAULT, object_enc_num_layers=2, camera_enc_num_layers=2, enc_nhead=2,
             diffusion_attn_num_heads=4, diffusion_num_layers=2)

# diffusion_model_type='default' (AMDDiffusionTransformerModelTempMotion): object stream only
SMALL_OBJECT_ONLY = dict(SMALL, use_camera=False, diffusion_model_type="default")

CONFIGS = {"default": DEFAULT, "small": SMALL, "small_object_only": SMALL_OBJECT_ONLY}


# the older AMDModel (model_AMD.py:33-594), two runnable small configurations
AMDMODEL_TEMPORAL = dict(video_frames=4, use_filter=True, use_grey=True, use_motiontemporal=True, use_mask=True, motion_type="decouple",
                         object_motion_token_num=4, object_motion_token_channel=16, object_enc_num_layers=2, enc_nhead=2, enc_ndim=64,
                         camera_motion_token_num=4, camera_motion_token_channel=8, camera_enc_num_layers=2, motion_token_num=4,
                         motion_token_channel=16, diffusion_model_type="spatial", diffusion_attn_num_heads=2, diffusion_num_layers=7)
AMDMODEL_LEGACY = dict(video_frames=4, use_filter=True, use_grey=True, use_motiontemporal=False, use_camera_down=True, motion_type="plus",
                       object_motion_token_num=4, object_motion_token_channel=16, object_enc_num_layers=2, enc_nhead=2, enc_ndim=64,
                       camera_motion_token_num=4, camera_motion_token_channel=16, camera_enc_num_layers=2, motion_token_num=4,
                       motion_token_channel=16, diffusion_model_type="default", diffusion_attn_num_heads=2, diffusion_num_layers=3)

# AMDModel_Camera (model_AMD.py:894-1661), runnable configuration: camera stream only
AMDMODEL_CAMERA = dict(video_frames=4, use_filter=True, filter_num=0.4, use_grey=True, use_mask=True, use_motiontemporal=False,
                       motion_type="decouple", camera_motion_token_num=4, camera_motion_token_channel=16, camera_enc_num_layers=2,
                       enc_nhead=2, enc_ndim=64, motion_token_num=4, motion_token_channel=16, diffusion_model_type="spatial",
                       diffusion_attn_num_heads=2, diffusion_num_layers=3)
# AMDModel_Rec (model_AMD.py:2840-3125): the non-diffusion reconstruction model
AMDMODEL_REC = dict(video_frames=4, motion_token_num=4, motion_token_channel=128, enc_num_layers=2, enc_nhead=2, enc_ndim=64,
                    motion_need_norm_out=True, diffusion_attn_num_heads=2, diffusion_num_layers=2)
