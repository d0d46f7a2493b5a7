"""Which kernel sits outside the north_star bound, and by how much?  ("What's weak" 1.)

For the parity cases in which the repo's attention kernels use the tolerance tail of tests/_cases.py::assert_close
(profiles/r02/parity_tail.json: a few elements per tensor beyond 2e-3 + 1e-2*|e|), run BOTH this library and the
flash-attn 2.8.3 wheel (the library the reference calls) on the SAME seeded inputs and count, against the same fp32 CPU
oracle, the elements each leaves outside the bound and the worst error / bound.  GPU box only; the oracle is used here
as the checker (tools/ is test infrastructure, not the product path).  Writes gpurun_out/tail_vs_flash_attn.json."""
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [os.path.join(ROOT, "micro-vllm_b200"), ROOT, os.path.join(ROOT, "tests")]
import torch  # noqa: E402
import b200ops  # noqa: E402
from flash_attn import flash_attn_with_kvcache as fa  # noqa: E402
from oracle.paged_attention import paged_attention_oracle  # noqa: E402
from _cases import make_paged_case, to_dev  # noqa: E402


def count(out, want):
    a, e = out.float().cpu(), want.float()
    err = (a - e).abs()
    lim = 2e-3 + 1e-2 * e.abs()
    return {"out_of_bound": int((err > lim).sum()), "worst_err_over_bound": round(float((err / lim).max()), 3),
            "max_abs_err": round(float(err.max()), 6), "mean_abs_err": float(err.mean()),
            "rms_err": float((err.double() ** 2).mean().sqrt())}


def decode_case(name, B, Hq, Hkv, D, lens, seed, splits):
    case = make_paged_case(B, Hq, Hkv, D, lens, seed=seed)
    want = paged_attention_oracle(case["q"], case["k_cache"].clone(), case["v_cache"].clone(), case["k"], case["v"],
                                  case["cos"], case["sin"], case["seqlens"], case["table"], case["scale"], True, False)[:, 0]
    rec = {"case": name, "numel": want.numel(), "kind": "decode"}
    for s in splits:
        d = to_dev(case)
        o = b200ops.paged_attn_decode(d["q"][:, 0], d["k_cache"], d["v_cache"], d["k"][:, 0], d["v"][:, 0], d["cos"], d["sin"],
                                      d["seqlens"], d["table"], softmax_scale=case["scale"], num_splits=s)
        torch.cuda.synchronize()
        rec[f"b200_splits{s}"] = count(o, want)
    for s in (1, 0):
        d = to_dev(case)
        o = fa(d["q"], d["k_cache"], d["v_cache"], k=d["k"], v=d["v"], rotary_cos=d["cos"], rotary_sin=d["sin"],
               cache_seqlens=d["seqlens"], block_table=d["table"], softmax_scale=case["scale"], causal=True,
               rotary_interleaved=False, num_splits=s)
        torch.cuda.synchronize()
        rec[f"flash_attn_splits{s}"] = count(o[:, 0], want)
    return rec


def prefill_case(name, B, S, Hq, Hkv, D, prefixes, seed):
    case = make_paged_case(B, Hq, Hkv, D, prefixes, S_new=S, seed=seed)
    want = paged_attention_oracle(case["q"], case["k_cache"].clone(), case["v_cache"].clone(), case["k"], case["v"],
                                  case["cos"], case["sin"], case["seqlens"], case["table"], case["scale"], True, False)
    rec = {"case": name, "numel": want.numel(), "kind": "prefill"}
    d = to_dev(case)
    o = b200ops.paged_attn_prefill(d["q"], d["k"], d["v"], d["k_cache"], d["v_cache"], d["cos"], d["sin"], d["seqlens"],
                                   d["table"], softmax_scale=case["scale"])
    torch.cuda.synchronize()
    rec["b200"] = count(o, want)
    d = to_dev(case)
    o = fa(d["q"], d["k_cache"], d["v_cache"], k=d["k"], v=d["v"], rotary_cos=d["cos"], rotary_sin=d["sin"],
           cache_seqlens=d["seqlens"], block_table=d["table"], softmax_scale=case["scale"], causal=True,
           rotary_interleaved=False)
    torch.cuda.synchronize()
    rec["flash_attn"] = count(o, want)
    return rec


def main():
    out = []
    lens_cm = [4095, 3000, 257, 5, 0, 2047, 2048, 1023, 64, 1]
    for ns in (2, 4, 5, 8):       # tests/test_gpu_decode.py::test_cluster_merge_equals_workspace_merge[*-28-4-128]
        out.append(decode_case(f"cluster_merge[{ns}-28-4-128]", len(lens_cm), 28, 4, 128, lens_cm, ns * 7 + 28, [ns, 0]))
    lens_v = [700, 255, 256, 31, 4000, 1, 0, 513]
    for v in (1, 5):              # test_all_kernel_variants[128-...]: same inputs, default variant
        out.append(decode_case(f"all_kernel_variants seed {10 + v}", 8, 32, 8, 128, lens_v, 10 + v, [3, 0]))
    out.append(decode_case("cfg2-like b8 4k", 8, 32, 8, 128, [4095] * 8, 3, [0]))
    # tests/test_gpu_prefill.py: from_empty_cache[2-300-32-8-128] (seed S + Hq), two_tiles[2-300-32-8-128-prefixes6]
    out.append(prefill_case("prefill_from_empty_cache[2-300-32-8-128]", 2, 300, 32, 8, 128, [0, 0], 332))
    out.append(prefill_case("prefill_from_empty_cache[1-512-28-4-128]", 1, 512, 28, 4, 128, [0], 540))
    out.append(prefill_case("chunked prefix 300 + 200", 2, 200, 8, 2, 128, [300, 223], 500))
    for r in out:
        print(json.dumps(r))
    os.makedirs(os.path.join(ROOT, "gpurun_out"), exist_ok=True)
    json.dump(out, open(os.path.join(ROOT, "gpurun_out", "tail_vs_flash_attn.json"), "w"), indent=1)


if __name__ == "__main__":
    main()
