"""Drop-in mirror of the reference's ``avdiff`` package for the sampling hot path.

Same module names, callables, signatures, shapes, dtypes, ``state_dict`` keys and error behaviour as
mauruszach/multimodal_diffusion's ``avdiff`` (utils.ops, utils.schedule_utils, models.mmdt,
models.heads.noise_heads, models.adapters, models.tokenizers, models.schedules,
models.infer.sample_clip); what runs underneath is libavdiff_b200.so (sm_100a CUDA) through its C ABI.
"""
