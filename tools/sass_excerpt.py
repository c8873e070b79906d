"""SASS listings of the hot kernels for profiles/sass/: `cuobjdump -sass` of libspmb.so per kernel, encodings stripped,
with a mnemonic histogram on top (what the ptxas claims in DESIGN.md can be checked against).
    python tools/sass_excerpt.py [out_dir]"""
import collections, os, re, subprocess, sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
SO = os.path.join(ROOT, "libspm_b200", "lib", "libspmb.so")
OUT = sys.argv[1] if len(sys.argv) > 1 else os.path.join(ROOT, "profiles", "sass")
KERNELS = {
    "exact_kernel_plain": "_Z12exact_kernelILi0ELb0EEv9ExactArgs",
    "myers_kernel_nw4_plain": "_Z12myers_kernelILi4ELi0ELb0EEv9MyersArgs",
    "myers_cut_kernel_nw4_plain": "_Z16myers_cut_kernelILi4ELi0ELb0EEv9MyersArgs",
    "pigeon_filter_kernel_wide_step4": "_Z20pigeon_filter_kernelILi0ELb1ELi4EEv16PigeonFilterArgs",
    "pigeon_verify_kernel_nw4_plain": "_Z20pigeon_verify_kernelILi4ELi0EEv16PigeonVerifyArgs",
    "journal_derive_kernel": "_Z21journal_derive_kernel17JournalDeriveArgs",
    "pack_chunks_kernel_chunksource": "_Z18pack_chunks_kernelI11ChunkSourceEvT_10SpmbLayoutP5uint2S4_",
    "myers_chain_verify_kernel_wpl1_plain": "_Z25myers_chain_verify_kernelILi1ELi0EEv9ChainArgs",
}
os.makedirs(OUT, exist_ok=True)
res = subprocess.run(["cuobjdump", "-res-usage", SO], capture_output=True, text=True).stdout
for name, sym in KERNELS.items():
    txt = subprocess.run(["cuobjdump", "-sass", "-fun", sym, SO], capture_output=True, text=True).stdout
    lines = []
    for l in txt.splitlines():
        m = re.match(r"\s*/\*([0-9a-f]{4,})\*/\s+(.*?);\s*/\* 0x[0-9a-f]+ \*/", l)
        if m:
            lines.append((m.group(1), m.group(2).strip()))
    hist = collections.Counter()
    for _, ins in lines:
        op = ins.split()[1] if ins.startswith("@") else ins.split()[0]
        hist[op.split(".")[0]] += 1
    usage = ""
    for i, l in enumerate(res.splitlines()):
        if sym in l:
            usage = res.splitlines()[i + 1].strip()
    with open(os.path.join(OUT, name + ".sass"), "w") as f:
        f.write(f"// {sym}\n// {usage}\n// {len(lines)} instructions; mnemonic histogram:\n")
        f.write("//   " + "  ".join(f"{k}:{v}" for k, v in hist.most_common()) + "\n")
        for a, ins in lines:
            f.write(f"/*{a}*/  {ins} ;\n")
    print(name, len(lines), usage[:60])
