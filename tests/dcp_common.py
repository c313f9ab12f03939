"""Shared scaffolding of the decode-context-parallel tests: builds, from one unsharded flash-layout problem, the token
shard that lives on a given rank (token t -> rank t % N at local position t // N, reference
vllm_metax/v1/attention/backends/flash_attn.py:421-431)."""
import torch

from oracle import torch_oracle as orc


def make_problem(S=4, H=8, Hkv=2, D=128, B=16, L=300, seed=31, dtype=torch.bfloat16):
    NB = S * ((L + B - 1) // B) + 2
    d = orc.make_flash_decode_inputs(S, H, Hkv, D, B, L, NB, dtype=dtype, seed=seed)
    d["seq_lens"][0] = 1      # only rank 0 holds a token
    d["seq_lens"][1] = L
    return d


def shard_for_rank(d, world, rank):
    """(key_cache, value_cache, block_table, local_seq_lens) of `rank`: its tokens packed densely into its own pages."""
    kc, vc = d["key_cache"], d["value_cache"]
    NBg, B, Hkv, D = kc.shape
    S = d["query"].shape[0]
    lens = d["seq_lens"].tolist()
    max_local = (max(lens) + world - 1) // world
    nblk = (max_local + B - 1) // B + 1
    lk = torch.zeros(S * nblk + 1, B, Hkv, D, dtype=kc.dtype)
    lv = torch.zeros_like(lk)
    table = torch.zeros(S, nblk, dtype=torch.int32)
    local_lens = []
    for s in range(S):
        blocks = [1 + s * nblk + j for j in range(nblk)]  # block 0 stays unused
        table[s] = torch.tensor(blocks, dtype=torch.int32)
        n = 0
        for t in range(rank, lens[s], world):
            gb, go = int(d["block_tables"][s, t // B]), t % B
            lk[blocks[n // B], n % B] = kc[gb, go]
            lv[blocks[n // B], n % B] = vc[gb, go]
            n += 1
        local_lens.append(n)
    return lk, lv, table, torch.tensor(local_lens, dtype=torch.int32)
