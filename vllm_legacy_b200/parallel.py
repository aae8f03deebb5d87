"""KV-head (tensor-parallel) sharding of the paged-attention path across GPUs.

The reference shards attention by KV head under TP (vllm/config.py:320-333): rank r owns
`num_kv_heads / tp` KV heads, the `num_heads / tp` query heads that read them, and its own
KV-cache shard `[num_blocks, num_kv_heads / tp, ...]`; block tables, seq_lens and slot mappings are
replicated (vllm/worker/worker.py:262-268).  Attention is independent per (sequence, KV head), so
the kernel needs NO collective: each process runs the same ops on its shard.
"""
from dataclasses import dataclass


@dataclass(frozen=True)
class HeadShard:
    rank: int
    world_size: int
    num_heads: int          # query heads on this rank
    num_kv_heads: int       # KV heads on this rank
    q_head_start: int       # first global query head
    kv_head_start: int      # first global KV head
    kv_replicas: int        # >1 when world_size > total KV heads (heads replicated, config.py:327)


def shard_heads(total_num_heads: int, total_num_kv_heads: int, world_size: int, rank: int) -> HeadShard:
    """Per-rank head counts exactly as ModelConfig.get_num_kv_heads / get_num_attention_heads
    (vllm/config.py:320-333): query heads divide evenly; KV heads are divided, or replicated when
    there are fewer KV heads than ranks (at least one per rank)."""
    if not (0 <= rank < world_size):
        raise ValueError(f"rank {rank} outside world of {world_size}")
    if total_num_heads % world_size != 0:
        raise ValueError(f"Total number of attention heads ({total_num_heads}) must be divisible by "
                         f"tensor parallel size ({world_size}).")
    if total_num_heads % total_num_kv_heads != 0:
        raise ValueError("num_heads must be a multiple of num_kv_heads")
    q_per_rank = total_num_heads // world_size
    if total_num_kv_heads >= world_size:
        if total_num_kv_heads % world_size != 0:
            raise ValueError("num_kv_heads must be divisible by the tensor parallel size")
        kv_per_rank = total_num_kv_heads // world_size
        return HeadShard(rank, world_size, q_per_rank, kv_per_rank, rank * q_per_rank, rank * kv_per_rank, 1)
    if world_size % total_num_kv_heads != 0:
        raise ValueError("tensor parallel size must be a multiple of num_kv_heads when it is larger")
    replicas = world_size // total_num_kv_heads
    return HeadShard(rank, world_size, q_per_rank, 1, rank * q_per_rank, rank // replicas, replicas)
