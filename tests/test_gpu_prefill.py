"""Paged prefill attention (fused RoPE + page scatter + tcgen05/TMEM causal attention) against
the fp32 oracle, through the reference-named operator surface."""
import pytest
import torch

import b200ops
from core.paged_attention import flash_attn_with_kvcache
from oracle.elementwise import bf16_ulp_diff
from oracle.paged_attention import paged_attention_oracle
from oracle.rotary import rope_eager_oracle
from _cases import assert_close, make_paged_case, to_dev

pytestmark = pytest.mark.gpu


def check_prefill(case, interleaved=False, causal=True, rotary=True, variant=0):
    d = to_dev(case)
    kc_ref, vc_ref = case["k_cache"].clone(), case["v_cache"].clone()
    want = paged_attention_oracle(case["q"], kc_ref, vc_ref, case["k"], case["v"],
                                  case["cos"] if rotary else None, case["sin"] if rotary else None,
                                  case["seqlens"], case["table"], case["scale"], causal=causal, interleaved=interleaved)
    got = b200ops.paged_attn_prefill(d["q"], d["k"], d["v"], d["k_cache"], d["v_cache"],
                                     d["cos"] if rotary else None, d["sin"] if rotary else None,
                                     d["seqlens"], d["table"], softmax_scale=case["scale"], causal=causal,
                                     rotary_interleaved=interleaved, variant=variant)
    torch.cuda.synchronize()
    assert torch.equal(d["v_cache"].cpu(), vc_ref), "V pages must be bit-exact"
    kd = d["k_cache"].cpu()
    assert int(bf16_ulp_diff(kd, kc_ref).max()) <= 1, "rotated K pages differ by more than 1 bf16 ulp"
    assert_close(got, want, what="prefill attention output")
    return got


@pytest.mark.parametrize("B,S,Hq,Hkv,D", [(1, 128, 4, 4, 128), (2, 128, 8, 2, 128), (1, 256, 4, 1, 128),
                                           (2, 100, 4, 2, 128), (1, 1, 4, 4, 128), (3, 129, 14, 2, 64),
                                           (1, 512, 28, 4, 128), (2, 300, 32, 8, 128)])
def test_prefill_from_empty_cache(B, S, Hq, Hkv, D):
    """The reference's prefill: cache_seqlens zeroed (model_prefill.py:32), positions 0..S-1."""
    case = make_paged_case(B, Hq, Hkv, D, [0] * B, S_new=S, seed=S + Hq)
    check_prefill(case)


@pytest.mark.parametrize("prefix,S", [(1, 127), (255, 2), (256, 128), (300, 200), (1000, 64), (129, 513)])
def test_chunked_prefill_on_cached_prefix(prefix, S):
    """configs[2]: a chunk of S new tokens on top of a cached prefix (bottom-right causal mask)."""
    case = make_paged_case(2, 8, 2, 128, [prefix, max(0, prefix - 77)], S_new=S, seed=prefix + S)
    check_prefill(case)


def test_prefill_long_prompt_many_tiles():
    """16 KV tiles per query tile: exercises the K/V ring phases, S double buffering and the lazy
    O rescale far beyond the first wrap-around."""
    case = make_paged_case(1, 8, 2, 128, [0], S_new=2048, seed=21)
    check_prefill(case)
    case = make_paged_case(1, 4, 4, 64, [1100], S_new=1500, seed=22)
    check_prefill(case)


def test_prefill_retired_variant_is_refused():
    """variant 1 (the first kernel: Q and P staged in shared memory) was retired in round 2: asking for it must fail
    loudly, not fall through to another kernel."""
    case = make_paged_case(1, 8, 2, 128, [0], S_new=64, seed=31)
    d = to_dev(case)
    with pytest.raises(b200ops.B200Error):
        b200ops.paged_attn_prefill(d["q"], d["k"], d["v"], d["k_cache"], d["v_cache"], d["cos"], d["sin"], d["seqlens"],
                                   d["table"], softmax_scale=case["scale"], variant=1)


@pytest.mark.parametrize("variant", [2, 3, 4, 5])
@pytest.mark.parametrize("D,Hq,Hkv", [(128, 8, 2), (64, 14, 2)])
def test_prefill_softmax_arithmetic_variants(variant, D, Hq, Hkv):
    """variants 2..5 = the TMEM-operand kernel with each softmax arithmetic (scalar; packed fp32x2;
    packed + 25 % / 50 % of the exponentials on the FMA-pipe polynomial): same bound against the oracle,
    exact zeros for masked keys (ragged prefixes, partial tiles, > 2 KV ring wraps)."""
    check_prefill(make_paged_case(2, Hq, Hkv, D, [0, 300], S_new=390, seed=41), variant=variant)
    check_prefill(make_paged_case(1, Hq, Hkv, D, [777], S_new=1100, seed=42), variant=variant)
    check_prefill(make_paged_case(2, Hq, Hkv, D, [5, 130], S_new=140, seed=43), causal=False, variant=variant)


PAIR_CASES = [  # B, S, Hq, Hkv, D, prefixes
    (1, 128, 4, 4, 128, [0]),            # G = 1 (odd): two consecutive row tiles per CTA, second one dead
    (2, 128, 8, 2, 128, [0, 0]),         # G = 4 (even): two heads per CTA
    (2, 100, 4, 2, 128, [0, 57]),        # partial tile
    (1, 1, 4, 4, 128, [0]),
    (3, 129, 14, 2, 64, [0, 1, 255]),    # G = 7, D = 64, odd number of row tiles
    (1, 512, 28, 4, 128, [0]),           # Qwen2.5-7B heads, 4 row tiles -> 2 CTAs per head
    (2, 300, 32, 8, 128, [0, 300]),      # Llama-3-8B heads on a cached prefix
    (1, 1100, 8, 2, 128, [777]),         # > 2 K/V ring wraps, ragged prefix
    (2, 390, 14, 2, 64, [0, 300]),
    (1, 700, 6, 6, 128, [131]),          # MHA, odd tile count, prefix not tile aligned
]


@pytest.mark.parametrize("variant", [6, 7, 8])
@pytest.mark.parametrize("B,S,Hq,Hkv,D,prefixes", PAIR_CASES)
def test_prefill_two_tiles_per_cta(variant, B, S, Hq, Hkv, D, prefixes):
    """variants 6..8 = two 128-row query tiles per CTA sharing each K/V tile (ping-pong), probabilities
    written over the scores in tensor memory; head pairs when G is even, row-tile pairs otherwise."""
    check_prefill(make_paged_case(B, Hq, Hkv, D, prefixes, S_new=S, seed=S + Hq + variant), variant=variant)


@pytest.mark.parametrize("variant", [6, 8])
def test_prefill_two_tiles_long_and_noncausal(variant):
    check_prefill(make_paged_case(1, 8, 2, 128, [0], S_new=2048, seed=51), variant=variant)
    check_prefill(make_paged_case(1, 7, 1, 128, [300], S_new=1500, seed=52), variant=variant)
    check_prefill(make_paged_case(2, 8, 4, 128, [5, 130], S_new=140, seed=53), causal=False, variant=variant)
    check_prefill(make_paged_case(2, 3, 3, 64, [5, 130], S_new=260, seed=54), causal=False, variant=variant)


PERSIST_CASES = [  # B, S, Hq, Hkv, D, prefixes, causal   (work items of the persistent kernel >> 148 CTAs)
    (3, 700, 32, 8, 128, [0, 131, 300], True),               # 288 items, 6..8 KV tiles, ragged prefixes
    (40, 100, 8, 2, 128, [(i * 37) % 90 for i in range(40)], True),   # 160 items of ONE or TWO KV tiles each
    (6, 520, 14, 2, 64, [0, 1, 255, 256, 77, 600], True),    # G = 7: row-tile pairs, dead second tile, D = 64
    (5, 260, 32, 8, 128, [5, 130, 0, 256, 300], False),      # non-causal, partial last tile (V tail zeroing)
    (2, 1500, 32, 8, 128, [0, 40], True),                    # 384 items, up to 13 KV tiles: several items per CTA
]


@pytest.mark.parametrize("variant", [0, 9, 10, 11])
@pytest.mark.parametrize("B,S,Hq,Hkv,D,prefixes,causal", PERSIST_CASES)
def test_prefill_persistent_many_items(variant, B, S, Hq, Hkv, D, prefixes, causal):
    """variants 9..11 (and the default, which selects 9 for these sizes) = ONE resident CTA per SM pulling
    (query-tile pair) work items from a device counter: every mbarrier phase keeps counting across items, the next
    item's Q / K / V loads and first Q.K^T overlap the previous item's last softmax and epilogue."""
    if variant in (10, 11) and S > 700:
        pytest.skip("softmax-arithmetic variants are covered on the smaller cases")
    check_prefill(make_paged_case(B, Hq, Hkv, D, prefixes, S_new=S, seed=S + Hq + variant), causal=causal,
                  variant=variant)


@pytest.mark.parametrize("B,S,Hq,Hkv,D,prefixes", PAIR_CASES)
def test_prefill_persistent_few_items(B, S, Hq, Hkv, D, prefixes):
    """Fewer items than SMs: every CTA runs one item and then meets the end-of-work sentinel."""
    check_prefill(make_paged_case(B, Hq, Hkv, D, prefixes, S_new=S, seed=S + Hq + 9), variant=9)


@pytest.mark.parametrize("B,S,Hq,Hkv,D,prefix", [(1, 8192, 32, 8, 128, 0), (4, 2048, 14, 2, 64, 0), (2, 2048, 28, 4, 128, 6144)])
def test_prefill_persistent_equals_one_cta_per_item_kernel_at_bench_sizes(B, S, Hq, Hkv, D, prefix):
    """At the sizes the TFLOP/s figures are quoted on (64 KV tiles per item, > 1000 items, hundreds of ring wraps per
    resident CTA) the fp32 CPU oracle is too slow to be the checker; the round-1 one-CTA-per-item kernel — itself checked
    against the oracle on every smaller case — is.  With the MUFU-only softmax arithmetic (variants 10 and 7) both kernels
    do the same operations in the same order per query tile, so the outputs must be BIT-IDENTICAL, whatever CTA ran an
    item and however the items were interleaved; so must the K / V pages."""
    page = 256
    per_seq = -(-(prefix + S) // page)
    n_blocks = B * per_seq + 3
    g = torch.Generator(device="cuda").manual_seed(S + Hq)
    dev = "cuda"
    table = torch.randperm(n_blocks, device=dev)[: B * per_seq].view(B, per_seq).to(torch.int32)
    lens = torch.full((B,), prefix, dtype=torch.int32, device=dev)
    q, k, v = (torch.randn(B, S, h, D, device=dev, generator=g).to(torch.bfloat16) for h in (Hq, Hkv, Hkv))
    from core.paged_attention import build_rotary_tables
    cos, sin = build_rotary_tables(D, prefix + S + 1)
    cos, sin = cos[:, : D // 2].contiguous().to(dev), sin[:, : D // 2].contiguous().to(dev)
    outs, pages = [], []
    for variant in (10, 7):
        kc = torch.randn(n_blocks, page, Hkv, D, device=dev, generator=torch.Generator(device="cuda").manual_seed(5)).to(torch.bfloat16)
        vc = torch.randn(n_blocks, page, Hkv, D, device=dev, generator=torch.Generator(device="cuda").manual_seed(6)).to(torch.bfloat16)
        outs.append(b200ops.paged_attn_prefill(q, k, v, kc, vc, cos, sin, lens, table, variant=variant))
        pages.append((kc, vc))
    torch.cuda.synchronize()
    assert torch.isfinite(outs[0]).all()
    assert torch.equal(outs[0], outs[1]), "persistent kernel and one-CTA-per-item kernel disagree"
    assert torch.equal(pages[0][0], pages[1][0]) and torch.equal(pages[0][1], pages[1][1])
    # and the default arithmetic (25 % polynomial exp2) stays within bf16 noise of it
    dflt = b200ops.paged_attn_prefill(q, k, v, pages[0][0], pages[0][1], cos, sin, lens, table)
    assert float((dflt.float() - outs[0].float()).abs().max()) < 2e-2


def test_prefill_persistent_counter_rearmed_every_call():
    """The work-item counter lives in the call's workspace and is zeroed by the RoPE/append kernel of the SAME call:
    back-to-back calls and CUDA-graph replays must each see a fresh counter (a stale one would skip all work)."""
    case = make_paged_case(3, 32, 8, 128, [0, 131, 300], S_new=700, seed=77)
    first = check_prefill(case)
    d = to_dev(case)
    args = (d["q"], d["k"], d["v"], d["k_cache"], d["v_cache"], d["cos"], d["sin"], d["seqlens"], d["table"])
    out = torch.empty_like(d["q"])
    for _ in range(3):
        out.zero_()
        b200ops.paged_attn_prefill(*args, softmax_scale=case["scale"], out=out)
        assert torch.equal(out, first)
    s = torch.cuda.Stream()
    s.wait_stream(torch.cuda.current_stream())
    with torch.cuda.stream(s):
        b200ops.paged_attn_prefill(*args, softmax_scale=case["scale"], out=out)
    torch.cuda.current_stream().wait_stream(s)
    g = torch.cuda.CUDAGraph()
    with torch.cuda.graph(g):
        b200ops.paged_attn_prefill(*args, softmax_scale=case["scale"], out=out)
    for _ in range(3):
        out.zero_()
        g.replay()
        torch.cuda.synchronize()
        assert torch.equal(out, first)


def test_prefill_interleaved_and_noncausal():
    case = make_paged_case(2, 8, 4, 128, [0, 130], S_new=140, seed=3)
    check_prefill(case, interleaved=True)
    case = make_paged_case(2, 8, 4, 128, [0, 130], S_new=140, seed=4)
    check_prefill(case, causal=False)


def test_reference_call_site_prerotated_inputs():
    """model_prefill.py:45-52: q/k rotated by the eager bf16 RoPE first, then
    flash_attn_with_kvcache(q, k_cache, v_cache, k=k, v=v, cache_seqlens=0, block_table, causal=True)
    with NO rotary tables; q/k/v are strided views of one qkv tensor."""
    B, S, H, D = 2, 192, 8, 128
    case = make_paged_case(B, H, H, D, [0, 0], S_new=S, seed=5)
    qkv = torch.stack((case["q"], case["k"], case["v"]), dim=2)            # [B, S, 3, H, D]
    q, k, v = qkv.unbind(dim=2)
    q_rot, k_rot = rope_eager_oracle(q, k, case["cos"], case["sin"])
    kc_ref, vc_ref = case["k_cache"].clone(), case["v_cache"].clone()
    want = paged_attention_oracle(q_rot, kc_ref, vc_ref, k_rot, v, None, None, case["seqlens"], case["table"],
                                  None, True, False)
    d = to_dev(case)
    qkv_d = qkv.cuda()
    qd, kd, vd = qkv_d.unbind(dim=2)
    from core.layer.rope import RoPE
    qd, kd = RoPE().forward(qd, kd, d["cos"], d["sin"])
    assert torch.equal(qd.cpu(), q_rot) or int(bf16_ulp_diff(qd.cpu(), q_rot).max()) <= 1
    out = flash_attn_with_kvcache(q=qd, k_cache=d["k_cache"], v_cache=d["v_cache"], k=kd, v=vd,
                                  cache_seqlens=d["seqlens"], block_table=d["table"], causal=True)
    assert out.shape == (B, S, H, D)
    assert_close(out, want)
    assert torch.equal(d["v_cache"].cpu(), vc_ref)


def test_prefill_then_decode_consistency():
    """encode -> decode round trip: decoding token S on top of a prefilled cache equals the last
    row of prefilling S + 1 tokens at once (same K/V bytes in the pages, same attention)."""
    B, S, Hq, Hkv, D = 2, 200, 32, 8, 128
    case = make_paged_case(B, Hq, Hkv, D, [0, 0], S_new=S + 1, seed=6, fill="zeros")
    d = to_dev(case)
    full = b200ops.paged_attn_prefill(d["q"], d["k"], d["v"], d["k_cache"], d["v_cache"], d["cos"], d["sin"],
                                      d["seqlens"], d["table"], softmax_scale=case["scale"])
    pages_k, pages_v = d["k_cache"].clone(), d["v_cache"].clone()
    d2 = to_dev(case)
    b200ops.paged_attn_prefill(d2["q"][:, :S].contiguous(), d2["k"][:, :S].contiguous(), d2["v"][:, :S].contiguous(),
                               d2["k_cache"], d2["v_cache"], d2["cos"], d2["sin"], d2["seqlens"], d2["table"],
                               softmax_scale=case["scale"])
    lens = torch.full((B,), S, dtype=torch.int32, device="cuda")
    last = b200ops.paged_attn_decode(d2["q"][:, S], d2["k_cache"], d2["v_cache"], d2["k"][:, S], d2["v"][:, S],
                                     d2["cos"], d2["sin"], lens, d2["table"], softmax_scale=case["scale"])
    assert torch.equal(d2["v_cache"], pages_v), "prefill(S) + decode must leave the same V pages as prefill(S+1)"
    assert torch.equal(d2["k_cache"], pages_k), "rotated K written by the decode kernel and the prefill kernel must be bit-identical"
    assert_close(last, full[:, S].float(), what="decode vs last prefill row")


@pytest.mark.parametrize("variant", [0, 2, 6, 9])
@pytest.mark.parametrize("D,Hq,Hkv", [(128, 8, 2), (64, 14, 2)])
def test_prefill_poisoned_cache_tail(variant, D, Hq, Hkv):
    """never-written cache slots may hold NaN / Inf (torch.empty caches, recycled pages).  Keys at or
    beyond seqlen_k are masked (P = 0), but the TMA box still brings their V rows into the tile, and 0 * NaN would
    poison every output row of the tile.  The kernel must zero those rows (flash-attn clears out-of-bounds V; the
    decode kernel zero-fills them): outputs of a NaN-filled pool must equal the oracle's, which never gathers them."""
    case = make_paged_case(2, Hq, Hkv, D, [0, 300], S_new=390, seed=61)
    case["k_cache"].fill_(float("nan"))
    case["v_cache"].fill_(float("nan"))
    # the cached prefix of row 1 (300 tokens) must be finite: that IS attended
    page = case["page"]
    g = torch.Generator().manual_seed(62)
    for t in range(300):
        blk, off = int(case["table"][1, t // page]), t % page
        case["k_cache"][blk, off] = torch.randn(Hkv, D, generator=g).to(torch.bfloat16)
        case["v_cache"][blk, off] = torch.randn(Hkv, D, generator=g).to(torch.bfloat16)
    d = to_dev(case)
    kc_ref, vc_ref = case["k_cache"].clone(), case["v_cache"].clone()
    want = paged_attention_oracle(case["q"], kc_ref, vc_ref, case["k"], case["v"], case["cos"], case["sin"],
                                  case["seqlens"], case["table"], case["scale"], causal=True, interleaved=False)
    assert torch.isfinite(want).all()
    got = b200ops.paged_attn_prefill(d["q"], d["k"], d["v"], d["k_cache"], d["v_cache"], d["cos"], d["sin"],
                                     d["seqlens"], d["table"], softmax_scale=case["scale"], variant=variant)
    torch.cuda.synchronize()
    assert torch.isfinite(got).all(), "NaN from never-written cache slots leaked into the output"
    assert_close(got, want, what="prefill attention output (poisoned pool)")
