"""Write compact SASS listings (instruction text only, no encodings) of the kernels the flagship steps launch into
profiles/sass/: one file per kernel instantiation.  `cuobjdump -sass` of the whole library is 34 MB; these are the ones a
training step, a sharded step and a decode step actually run."""
import os, re, subprocess, sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
SO = os.path.join(ROOT, "pretraining_llm_b200", "ops", "libb200llm.so")
OUT = os.path.join(ROOT, "profiles", "sass")
WANT = [
    r"gemm2_kernel<256, false, false, false, false>", r"gemm2_kernel<256, false, true, false, false>",
    r"gemm2_kernel<256, true, true, true, false>", r"gemm2_kernel<128, true, true, true, false>",
    r"gemm2_kernel<256, false, false, false, true>", r"gemm_kernel<128, false, false, false>",
    r"gemm_skinny_kernel<32>", r"attn_fwd_tc_kernel<64>", r"attn_fwd2_tc_kernel<128>",
    r"attn_bwd_tc_kernel<64, true>", r"attn_bwd_tc_kernel<64, false>", r"attn_bwd_tc_kernel<128, true>",
    r"attn_bwd_tc_kernel<128, false>", r"attn_bwd_delta_tc_kernel<64>", r"ln_fwd_kernel<3, 2>", r"ln_bwd_kernel<3, 2, true>",
    r"ln_bwd_finalize2_kernel", r"adamw_kernel", r"ce_grad_kernel", r"ce_finalize_kernel", r"embed_fwd_kernel",
    r"embed_bwd_tok_kernel", r"embed_bwd_tok_sorted_kernel", r"colsum_bf16_kernel",
    r"dp_reduce_adamw_kernel<8, true, 4>", r"dp_reduce_adamw_kernel<8, false, 1>", r"zero_reduce_adamw_kernel<8, true, 4>",
    r"zero_reduce_adamw_kernel<8, false, 1>", r"zero_wait_kernel", r"embed_fwd_sharded_kernel",
    r"gemm2_kernel<256, false, true, false, true>", r"vp_ce_finalize_kernel",
    r"vp_gather_rows_kernel", r"attn_decode_kernel<128>", r"sample_kernel",
]


def main():
    os.makedirs(OUT, exist_ok=True)
    for f in os.listdir(OUT):
        os.remove(os.path.join(OUT, f))
    txt = subprocess.run(["cuobjdump", "-sass", SO], capture_output=True, text=True).stdout
    blocks = re.split(r"\n\s*Function : ", txt)
    index = []
    for blk in blocks[1:]:
        mangled = blk.split("\n", 1)[0].strip()
        name = subprocess.run(["c++filt", mangled], capture_output=True, text=True).stdout.strip()
        short = re.sub(r"\(.*", "", name.replace("(anonymous namespace)::", "")).replace("void ", "")
        if not any(w == short or short.endswith(w) for w in WANT):
            continue
        lines = []
        for ln in blk.split("\n")[1:]:
            m = re.match(r"\s+/\*([0-9a-f]{4,})\*/\s+(.*?);\s*/\*", ln)
            if m:
                lines.append(f"{m.group(1)}  {m.group(2).strip()}")
        fn = re.sub(r"[^A-Za-z0-9_]+", "_", short).strip("_") + ".sass"
        with open(os.path.join(OUT, fn), "w") as f:
            f.write(f"// {name}\n// {len(lines)} instructions, sm_100a, from `cuobjdump -sass libb200llm.so` (tools/sass_listing.py)\n")
            f.write("\n".join(lines) + "\n")
        index.append((fn, len(lines)))
    with open(os.path.join(OUT, "README.md"), "w") as f:
        f.write("# SASS listings of the kernels the flagship steps launch (instruction text, encodings stripped)\n\n"
                "| file | instructions |\n|---|---|\n" + "".join(f"| `{a}` | {b} |\n" for a, b in sorted(index)))
    print(len(index), "kernels written")


if __name__ == "__main__":
    main()
