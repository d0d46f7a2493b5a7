"""Pin the elementwise oracle (oracle/elementwise.py) and the sm_100a kernels on the reference's OWN
Triton kernels (kernel/swiglu.py, kernel/rmsnorm_add.py, kernel/rmsnorm_residual.py), which only run
on a GPU (their wrappers assert ``is_cuda``; Triton's CPU interpreter truncates instead of rounding,
SURVEY.md §8c O1).

The reference tree is not part of this repo and does not exist on the GPU box, so the three kernel
files are staged — unmodified — into the git-ignored ``baseline/_ref/kernel/`` by
``tests/golden/stage_reference.sh`` in the build container; the staged copy travels with the gpurun
snapshot.  Then on a B200:

    gpurun -- python tests/golden/make_golden_triton.py     -> gpurun_out/triton_gpu.pt (+ .log)

and the file is committed as ``tests/golden/triton_gpu.pt``.  Inputs are regenerated from seeds by the
tests (sha256 of the input bytes is stored); outputs of the compiled Triton kernels are stored as bf16.
The script also prints, per case, the ulp distance of (a) the CPU oracle and (b) this repo's CUDA
kernels from the Triton outputs, and the time of both kernels on the same inputs."""
import hashlib
import importlib.util
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
REF_KERNEL = os.path.join(ROOT, "baseline", "_ref", "kernel")

SWIGLU_CASES = [(4, 14336, 11), (7, 4864, 12), (3, 100, 13), (16, 3584, 14), (2, 28672 // 2, 15)]
NORM_CASES = [(4, 4096, 21), (5, 896, 22), (3, 3584, 23), (4, 8192, 24), (2, 100, 25), (2, 16384, 26)]
EPS = 1e-6


def load_ref(name):
    path = os.path.join(REF_KERNEL, name + ".py")
    spec = importlib.util.spec_from_file_location("ref_kernel_" + name, path)
    mod = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(mod)
    return mod


def digest(*ts):
    h = hashlib.sha256()
    for t in ts:
        h.update(t.contiguous().view(torch.int16).numpy().tobytes())
    return h.hexdigest()


def swiglu_inputs(rows, d, seed):
    g = torch.Generator().manual_seed(seed)
    return (torch.randn(rows, 2 * d, generator=g) * 2).to(torch.bfloat16)


def norm_inputs(rows, n, seed):
    g = torch.Generator().manual_seed(seed)
    x = (torch.randn(rows, n, generator=g) * 3).to(torch.bfloat16)
    r = (torch.randn(rows, n, generator=g) * 2).to(torch.bfloat16)
    w = (1 + 0.2 * torch.randn(n, generator=g)).to(torch.bfloat16)
    return x, r, w


def _time(fn, reps=50):
    for _ in range(5):
        fn()
    torch.cuda.synchronize()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for _ in range(reps):
        fn()
    b.record()
    torch.cuda.synchronize()
    return a.elapsed_time(b) / reps * 1e3


def main():
    sys.path[:0] = [os.path.join(ROOT, "micro-vllm_b200"), ROOT]
    import b200ops
    from oracle.elementwise import bf16_ulp_diff, rmsnorm_oracle, rmsnorm_residual_oracle, swiglu_oracle
    ref_swiglu, ref_add, ref_res = load_ref("swiglu"), load_ref("rmsnorm_add"), load_ref("rmsnorm_residual")
    import triton
    gold = {"meta": {"torch": torch.__version__, "triton": triton.__version__,
                     "device": torch.cuda.get_device_name(0), "eps": EPS}}
    lines = []

    def stat(name, what, got, want):
        d = bf16_ulp_diff(got.cpu(), want.cpu())
        line = f"{name}: {what} vs Triton  max ulp {int(d.max())}  bit-exact {float((d == 0).float().mean()):.5f}"
        print(line)
        lines.append(line)

    for rows, d, seed in SWIGLU_CASES:
        gu = swiglu_inputs(rows, d, seed)
        dev = gu.cuda()
        t_fused = ref_swiglu.swiglu_fused(dev)
        buf = torch.empty(rows, d, dtype=torch.bfloat16, device="cuda")
        t_gemm = ref_swiglu.swiglu_gemm(dev, buf)
        torch.cuda.synchronize()
        name = f"swiglu_{rows}x{d}"
        assert torch.equal(t_fused, t_gemm), name + ": swiglu_fused != swiglu_gemm in the reference"
        ours = b200ops.swiglu(dev, torch.empty_like(buf))
        stat(name, "oracle", swiglu_oracle(gu), t_fused)
        stat(name, "b200  ", ours, t_fused)
        gold[name] = dict(kind="swiglu", rows=rows, d=d, seed=seed, digest=digest(gu), out=t_fused.cpu())

    for rows, n, seed in NORM_CASES:
        x, r, w = norm_inputs(rows, n, seed)
        xd, rd, wd = x.cuda(), r.cuda(), w.cuda()
        name = f"rmsnorm_{rows}x{n}"
        t_y = ref_add.rmsnorm(xd, wd, EPS)
        t_y2 = ref_add.rmsnorm_(xd, wd, torch.empty_like(xd), EPS)
        torch.cuda.synchronize()
        same = torch.equal(t_y, t_y2)
        print(f"{name}: reference rmsnorm == rmsnorm_ : {same}")
        lines.append(f"{name}: reference rmsnorm == rmsnorm_ : {same}")
        ours = b200ops.rmsnorm(xd, wd, torch.empty_like(xd), EPS)
        stat(name, "oracle", rmsnorm_oracle(x, w, EPS), t_y)
        stat(name, "b200  ", ours, t_y)
        stat(name + "_(gemm variant)", "b200  ", ours, t_y2)
        gold[name] = dict(kind="rmsnorm", rows=rows, n=n, seed=seed, digest=digest(x, r, w), y=t_y.cpu(),
                          y_gemm=t_y2.cpu())

        name = f"rmsnorm_residual_{rows}x{n}"
        a_y, a_res = ref_add.rmsnorm_residual_fused(xd, rd, wd, EPS)                # rmsnorm_add.py:89-113
        b_y, b_res = ref_res.rmsnorm_residual_fused(xd, rd, wd, EPS)                # rmsnorm_residual.py:44-68
        c_y, c_res = ref_res.rmsnorm_residual_gemm(xd, rd, wd, torch.empty_like(xd), torch.empty_like(xd), EPS)
        d_y, d_res = ref_res.rmsnorm_residual_hybrid(xd, rd, wd, torch.empty_like(xd), EPS)
        torch.cuda.synchronize()
        eq = [torch.equal(a_y, c_y), torch.equal(b_y, c_y), torch.equal(d_y, c_y),
              torch.equal(a_res, c_res), torch.equal(b_res, c_res), torch.equal(d_res, c_res)]
        line = f"{name}: reference variants equal to rmsnorm_residual_gemm (y: add.fused, res.fused, hybrid; res: same) {eq}"
        print(line)
        lines.append(line)
        o_y, o_res = b200ops.rmsnorm_residual(xd, rd, wd, torch.empty_like(xd), torch.empty_like(xd), EPS)
        w_y, w_res = rmsnorm_residual_oracle(x, r, w, EPS)
        stat(name + " y  ", "oracle", w_y, c_y)
        stat(name + " res", "oracle", w_res, c_res)
        stat(name + " y  ", "b200  ", o_y, c_y)
        stat(name + " res", "b200  ", o_res, c_res)
        gold[name] = dict(kind="rmsnorm_residual", rows=rows, n=n, seed=seed, digest=digest(x, r, w),
                          y=c_y.cpu(), res=c_res.cpu(), y_add_fused=a_y.cpu(), variants_equal=eq)

    # the kernels to beat, same inputs, decode-sized and prefill-sized (warm L2 for the small ones)
    for rows, d in [(32, 14336), (8192, 14336)]:
        gu = torch.randn(rows, 2 * d, device="cuda").to(torch.bfloat16)
        buf = torch.empty(rows, d, dtype=torch.bfloat16, device="cuda")
        t_ref = _time(lambda: ref_swiglu.swiglu_gemm(gu, buf))
        t_our = _time(lambda: b200ops.swiglu(gu, buf))
        line = f"time swiglu {rows}x{d}: Triton {t_ref:.2f} us, b200 {t_our:.2f} us"
        print(line); lines.append(line)
    for rows, n in [(32, 4096), (8192, 4096)]:
        x = torch.randn(rows, n, device="cuda").to(torch.bfloat16)
        r = torch.randn(rows, n, device="cuda").to(torch.bfloat16)
        w = torch.ones(n, device="cuda", dtype=torch.bfloat16)
        y, res = torch.empty_like(x), torch.empty_like(x)
        t_ref = _time(lambda: ref_res.rmsnorm_residual_gemm(x, r, w, y, res, EPS))
        t_our = _time(lambda: b200ops.rmsnorm_residual(x, r, w, y, res, EPS))
        line = f"time rmsnorm_residual {rows}x{n}: Triton {t_ref:.2f} us, b200 {t_our:.2f} us"
        print(line); lines.append(line)
        t_ref = _time(lambda: ref_add.rmsnorm_(x, w, y, EPS))
        t_our = _time(lambda: b200ops.rmsnorm(x, w, y, EPS))
        line = f"time rmsnorm {rows}x{n}: Triton {t_ref:.2f} us, b200 {t_our:.2f} us"
        print(line); lines.append(line)

    out_dir = os.path.join(ROOT, "gpurun_out")
    os.makedirs(out_dir, exist_ok=True)
    path = os.path.join(out_dir, "triton_gpu.pt")
    torch.save(gold, path)
    with open(os.path.join(out_dir, "triton_golden.log"), "w") as f:
        f.write("\n".join(lines) + "\n")
    print("wrote", path, os.path.getsize(path), "bytes")


if __name__ == "__main__":
    main()
