"""Instruction census of the steady-state loop of a kernel, from the SASS of an object file / cubin:

    python scripts/sass_loop_census.py lmc_atomi_b200/build/_lmc_b200/mix_kernels.o 'mix_chain_kernelI2f2Li1ELi0ELi5ELb1'

Prints, for the innermost loops (smallest bodies ending in a backward branch), the opcode histogram, the number of
packed FP32 instructions (FADD2 / FMUL2 / FFMA2, which hold the dispatch port for two cycles on B200 --
profiles/r02c_ubench_issue.txt) and the resulting dispatch-slot count.  Used to freeze
lmc_atomi_b200/roofline.py::MIX_PACKED_FRACTION."""
import collections
import os
import re
import subprocess
import sys
import tempfile


def cubin_of(path):
    if path.endswith(".cubin"):
        return path
    d = tempfile.mkdtemp()
    subprocess.run(["cuobjdump", "-xelf", "all", os.path.abspath(path)], cwd=d, check=True, capture_output=True)
    c = [f for f in os.listdir(d) if f.endswith(".cubin")]
    return os.path.join(d, c[0])


def loops(cubin, pattern):
    names = subprocess.run(["cuobjdump", "-sass", cubin], capture_output=True, text=True).stdout
    funs = [m.group(1) for m in re.finditer(r"Function : (\S+)", names) if pattern in m.group(1)]
    for fn in funs:
        out = subprocess.run(["cuobjdump", "-sass", "-fun", fn, cubin], capture_output=True, text=True).stdout
        ins = []
        for l in out.split("\n"):
            m = re.match(r"\s+/\*([0-9a-f]{4,6})\*/\s+(.*?);", l)
            if m:
                ins.append((int(m.group(1), 16), m.group(2).strip()))
        found = []
        for a, t in ins:
            m = re.search(r"\bBRA\b.*0x([0-9a-f]+)", t)
            if m and int(m.group(1), 16) < a:
                tgt = int(m.group(1), 16)
                body = [(b, x) for b, x in ins if tgt <= b <= a]
                # blocks skipped by a predicated FORWARD branch inside the loop (here: the shifted re-evaluation taken
                # only when every mixture component underflows) are reported apart from the steady-state path
                cold = set()
                for b, x in body:
                    f = re.search(r"^@!?U?P\d+\s+BRA\b.*0x([0-9a-f]+)", x)
                    if f and b < int(f.group(1), 16) <= a:
                        cold.update(c for c, _ in body if b < c < int(f.group(1), 16))
                found.append(([x for b, x in body if b not in cold], [x for b, x in body if b in cold]))
        yield fn, sorted(found, key=lambda hc: len(hc[0]) + len(hc[1]))


def census(body):
    cnt = collections.Counter()
    for t in body:
        op = t.split()[1] if t.startswith("@") else t.split()[0]
        cnt[op.split(".")[0]] += 1
    packed = cnt["FADD2"] + cnt["FMUL2"] + cnt["FFMA2"]
    return cnt, packed


if __name__ == "__main__":
    cubin = cubin_of(sys.argv[1])
    for fn, ls in loops(cubin, sys.argv[2]):
        print(fn)
        for body, cold in ls[:int(sys.argv[3]) if len(sys.argv) > 3 else 2]:
            cnt, packed = census(body)
            n = len(body)
            print(f"  loop: {n} instructions on the steady-state path (+{len(cold)} in forward-branch-skipped blocks): "
                  f"packed {packed} ({packed / n:.3f}), MUFU {cnt['MUFU']}, IMAD {cnt['IMAD']}, "
                  f"dispatch slots (packed x2) {n + packed}")
            print("   ", dict(cnt.most_common(16)))
