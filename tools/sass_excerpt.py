"""Regenerates profiles/r2_sass_excerpt.txt: static counts of the memory / synchronisation instructions of the hot
kernels in the built library (cuobjdump -sass), with one sample line each.  Runs without a GPU."""
import collections
import os
import re
import subprocess

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sass = subprocess.run(["cuobjdump", "-sass", os.path.join(ROOT, "paddle_ray_b200", "libpaddle_ray_b200.so")],
                      capture_output=True, text=True).stdout
funcs = re.split(r'\n\s*Function : ', sass)[1:]
want = {
    "adam_chunks_kernel<bf16, bf16, master>": "adam_chunks_kernelI13__nv_bfloat16S2_Lb1E",
    "adam_chunks_kernel<float, float, no master>": "adam_chunks_kernelIffLb0E",
    "grad_pass_kernel<bf16, mode 2 (pack), norm, no mp>": "grad_pass_kernelI13__nv_bfloat16Li2ELb1ELb0E",
    "map_chunks_kernel<float, ScaleOp>": "map_chunks_kernelIfNS0_7ScaleOp",
    "norm_finalize_kernel": "norm_finalize_kernel",
    "nvl_exchange_chunks_kernel<bf16, 8 ranks, no fold, norm>": "nvl_exchange_chunks_kernelI13__nv_bfloat16Li8ELb0ELb1ELb0E",
    "nvl_exchange_kernel<float, 8 ranks, no fold>": "nvl_exchange_kernelIfLi8ELb0E",
    "nvl_gather_chunks_kernel<bf16> (sharded step: parameter gather)": "nvl_gather_chunks_kernelI13__nv_bfloat16E",
    "sparse_adam_kernel<float, no master, 4 columns per lane>": "sparse_adam_kernelIfLb0ELi4E",
    "merge_rows_kernel<float, 4>": "merge_rows_kernelIfLi4E",
}
out = ["# SASS excerpt (round 2): cuobjdump -sass paddle_ray_b200/libpaddle_ray_b200.so, sm_100a cubins only",
       "# (tools/sass_excerpt.py).  Per kernel: static counts of the instructions that carry the design.",
       "#   LDG.E.NA.ENL2.256 / STG.E.ENL2.256 = 256-bit streaming global access (one fp32 lane vector)",
       "#   LDG.E.NA.128 / STG.E.128           = 128-bit (one lane vector of a 16-bit type)",
       "#   ACQBULK / PREEXIT                  = griddepcontrol.wait / .launch_dependents (programmatic dependent launch)",
       "#   LDG.E.STRONG.GPU + ATOMG           = coherent beta-pow loads and the ticket whose increment depends on them",
       "#   MEMBAR.ALL.SYS                     = fence.acq_rel.sys and the fence half of st.release.sys (exchange barriers)",
       "# No TMA / tcgen05 / MMA instruction anywhere: nothing on this path is a contraction.", ""]
pat = re.compile(r'\b(LDG\.E[A-Z0-9.]*|STG\.E[A-Z0-9.]*|ATOMG[A-Z0-9.]*|MEMBAR[A-Z.]*|ACQBULK|PREEXIT|SHFL[A-Z.]*|BAR\.[A-Z.]*'
                 r'|HMUL2[A-Z0-9.]*|F2FP[A-Z0-9._]*|MUFU[A-Z0-9.]*|UTMA[A-Z0-9.]*|UTC[A-Z0-9.]*|HMMA[A-Z0-9.]*)\b')
for title, key in want.items():
    for f in funcs:
        name = f.split('\n', 1)[0].strip()
        if key in name:
            c = collections.Counter(pat.findall(f))
            out.append("## %s\n#   %s\n#   %d instructions" % (title, name, len(re.findall(r'/\*[0-9a-f]{4}\*/', f))))
            out += ["    %5d  %s" % (v, k) for k, v in sorted(c.items(), key=lambda kv: (-kv[1], kv[0]))]
            for k in ("LDG.E.NA.ENL2.256", "LDG.E.NA.128", "STG.E.ENL2.256", "STG.E.128", "ACQBULK", "PREEXIT", "ATOMG", "MEMBAR.ALL.SYS"):
                m = re.search(r'/\*[0-9a-f]{4}\*/\s+([^\n;]*%s[^\n;]*;)' % re.escape(k), f)
                if m:
                    out.append("    e.g.  " + m.group(1).strip())
            out.append("")
            break
    else:
        out.append("## %s: (not found)\n" % title)
tot = collections.Counter(re.findall(r'\b(LDG\.E\.NA\.ENL2\.256|STG\.E\.ENL2\.256|LDG\.E\.NA\.128|STG\.E\.128)\b', sass))
out.append("## whole library: " + ", ".join("%s x %d" % kv for kv in sorted(tot.items())))
open(os.path.join(ROOT, "profiles", "r2_sass_excerpt.txt"), "w").write("\n".join(out) + "\n")
print("\n".join(out[-3:]))
