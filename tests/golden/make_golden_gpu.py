"""Generate attention golden vectors with the library the reference calls (flash-attn wheel)
on a B200:   gpurun -- python tests/golden/make_golden_gpu.py   -> gpurun_out/flash_attn_gpu.pt
(copy the file to tests/golden/).  Also pins the CPU oracle: prints the max deviation of
oracle/paged_attention.py from flash-attn on the same inputs."""
import hashlib
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path[:0] = [os.path.join(ROOT, "tests"), os.path.join(ROOT, "tests", "golden"), ROOT]
from flash_attn import flash_attn_with_kvcache  # noqa: E402
from _cases import make_paged_case  # noqa: E402
from oracle.paged_attention import paged_attention_oracle  # noqa: E402
from oracle.elementwise import bf16_ulp_diff  # noqa: E402
from make_golden_gpu_lib import appended_rows, inputs_digest  # noqa: E402


CASES = {
    # name: (make_paged_case kwargs, interleaved, rotary)
    "decode_neox_gqa4_d128": (dict(B=6, Hq=8, Hkv=2, D=128, lens=[0, 100, 255, 256, 300, 511], seed=1), False, True),
    "decode_interleaved_mha_d64": (dict(B=4, Hq=4, Hkv=4, D=64, lens=[1, 254, 257, 40], seed=2), True, True),
    "decode_gqa7_d64": (dict(B=3, Hq=14, Hkv=2, D=64, lens=[129, 0, 255], seed=3), False, True),
    "decode_llama8b_b4_4k": (dict(B=4, Hq=32, Hkv=8, D=128, lens=[4095, 4000, 2049, 1], seed=6), False, True),
    # round 2: BASELINE.json configs[1] at full size, one 128k-context row (configs[4]), D = 64 and G = 7 prefill
    "decode_llama8b_b32_4k_cfg2": (dict(B=32, Hq=32, Hkv=8, D=128, lens=[4095] * 32, seed=7), False, True),
    "decode_llama8b_b1_128k": (dict(B=1, Hq=32, Hkv=8, D=128, lens=[131071], seed=8, rot_rows=131080), False, True),
    "prefill_gqa7_d64": (dict(B=1, Hq=7, Hkv=1, D=64, lens=[0], S_new=300, seed=9), False, True),
    "prefill_gqa7_d128_prefix": (dict(B=1, Hq=7, Hkv=1, D=128, lens=[300], S_new=200, seed=10), False, True),
    "prefill_empty_d128": (dict(B=1, Hq=8, Hkv=2, D=128, lens=[0], S_new=150, seed=4), False, False),
    "prefill_prefix_d128": (dict(B=1, Hq=4, Hkv=1, D=128, lens=[200], S_new=130, seed=5), False, True),
}


def record(name, kwargs, interleaved, rotary, out_dict):
    case = make_paged_case(**kwargs)
    d = {k: (v.cuda() if torch.is_tensor(v) else v) for k, v in case.items()}
    kw = dict(rotary_cos=d["cos"], rotary_sin=d["sin"]) if rotary else {}
    out = flash_attn_with_kvcache(d["q"], d["k_cache"], d["v_cache"], k=d["k"], v=d["v"],
                                  cache_seqlens=d["seqlens"], block_table=d["table"], causal=True,
                                  rotary_interleaved=interleaved, **kw)
    torch.cuda.synchronize()
    kc, vc = case["k_cache"].clone(), case["v_cache"].clone()
    want = paged_attention_oracle(case["q"], kc, vc, case["k"], case["v"], case["cos"] if rotary else None,
                                  case["sin"] if rotary else None, case["seqlens"], case["table"], None, True, interleaved)
    err = (out.float().cpu() - want).abs()
    lim = 2e-3 + 1e-2 * want.abs()
    v_same = torch.equal(d["v_cache"].cpu(), vc)
    k_ulp = int(bf16_ulp_diff(d["k_cache"].cpu(), kc).max())
    print(f"{name}: oracle vs flash-attn  max|out err|={err.max().item():.3e}  out-of-tolerance={int((err > lim).sum())}"
          f"  V bit-exact={v_same}  K max ulp diff={k_ulp}")
    out_dict[name] = dict(kwargs=kwargs, interleaved=interleaved, rotary=rotary, digest=inputs_digest(case),
                          out=out.cpu(), k_rows=appended_rows(d["k_cache"].cpu(), case),
                          v_rows=appended_rows(d["v_cache"].cpu(), case), torch=torch.__version__)


def main():
    gold = {}
    for name, (kwargs, interleaved, rotary) in CASES.items():
        record(name, kwargs, interleaved, rotary, gold)
    os.makedirs(os.path.join(ROOT, "gpurun_out"), exist_ok=True)
    path = os.path.join(ROOT, "gpurun_out", "flash_attn_gpu.pt")
    torch.save(gold, path)
    print("wrote", path, os.path.getsize(path), "bytes")


if __name__ == "__main__":
    main()
