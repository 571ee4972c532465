"""mlx_parallm_b200 -- the batched decode step of mlx_parallm on hand-written sm_100a CUDA.

Public surface mirrors the reference package (mlx_parallm): ``utils.{load, generate_step, batch_generate,
generate, stream_generate, batch_stream_generate_text, batch_generate_text}``,
``models.{llama, qwen3}.{Model, ModelArgs}``, ``models.base.{BatchedKVCache, PagedKVCache}``,
``sample_utils.top_p_sampling``.  The arithmetic is in libb200parallm.so (include/b200p.h).
"""
__version__ = "0.1.0"
