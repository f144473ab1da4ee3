"""cuobjdump -sass of every kernel the pipeline launches at HEAD -> profiles/<tag>_sass/<kernel>.sass (one file per
instantiation that the default path uses), plus an instruction-mix summary (profiles/<tag>_sass/SUMMARY.md).

    python tools/dump_sass.py r2"""
import collections
import os
import re
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
LIB = os.path.join(ROOT, "shape_based_matching_b200", "libsbm_b200.so")
# the instantiations a default run launches (M1 / C4 / C5: colour and gray frames, T = {4, 8}) and the fallbacks
WANTED = [
    r"k_orient_mma<3, 5>", r"k_orient_mma<3, 3>", r"k_orient_mma<3, 2>", r"k_orient_mma<1, 5>",  # K1, default (36 / 20 / 12 rows)
    r"k_orient<3, 32, false>", r"k_orient<3, 16, false>", r"k_orient<1, 32, false>",  # K1 with SBM_ORIENT_MMA=0
    r"k_orient<3, 16, true>",  # training front end (magnitude / angle dumps)
    r"k_pyrdown<3>", r"k_pyrdown<1>",
    r"k_spread_linearize<4, false, true, false", r"k_spread_linearize<8, false, false, true", r"k_spread_linearize<8, true, false, false",
    r"k_similarity_bits<3, false>", r"k_similarity_bits<6, false>", r"k_similarity_bits<9, false>", r"k_similarity_bits<3, true>",
    r"k_similarity_nib", r"k_similarity\(", r"k_refine_tiled", r"k_refine\(", r"k_export", r"k_overflow_flag",
]


def main():
    tag = sys.argv[1] if len(sys.argv) > 1 else "head"
    out_dir = os.path.join(ROOT, "profiles", tag + "_sass")
    os.makedirs(out_dir, exist_ok=True)
    sass = subprocess.run(["cuobjdump", "-sass", LIB], stdout=subprocess.PIPE, text=True, check=True).stdout
    names = subprocess.run(["cu++filt"], input="\n".join(re.findall(r"Function : (\S+)", sass)), stdout=subprocess.PIPE, text=True).stdout.split("\n")
    blocks = re.split(r"(?=\t\tFunction : )", sass)[1:]
    commit = subprocess.run(["git", "-C", ROOT, "rev-parse", "--short", "HEAD"], stdout=subprocess.PIPE, text=True).stdout.strip()
    rows = []
    for blk, name in zip(blocks, names):
        name = name.replace("sbm::", "").replace("void ", "").replace("(int)", "").replace("(bool)0", "false").replace("(bool)1", "true")
        hit = [w for w in WANTED if re.search(w.replace("<", r"<").replace("(", r"\(") if not w.endswith(r"\(") else w, name)]
        if not hit:
            continue
        short = re.sub(r"\(.*", "", name)
        fname = re.sub(r"[^A-Za-z0-9_]+", "_", short).strip("_") + ".sass"
        ops = collections.Counter()
        n = 0
        for line in blk.split("\n"):
            m = re.match(r"\s+/\*[0-9a-f]{4,}\*/\s+(@!?U?P\d\s+)?([A-Z0-9_.]+)", line)
            if m:
                ops[m.group(2).split(".")[0]] += 1
                n += 1
        with open(os.path.join(out_dir, fname), "w") as f:
            f.write("// %s\n// cuobjdump -sass shape_based_matching_b200/libsbm_b200.so at commit %s (+ working tree)\n" % (name, commit))
            # instruction text only: the hex encodings (the trailing comment and the continuation line) triple the size
            for line in blk.split("\n"):
                if re.match(r"\s+/\* 0x[0-9a-f]+ \*/\s*$", line):
                    continue
                f.write(re.sub(r"\s*/\* 0x[0-9a-f]+ \*/\s*$", "", line) + "\n")
        rows.append((short, n, ops))
    with open(os.path.join(out_dir, "SUMMARY.md"), "w") as f:
        f.write("# SASS at %s: static instruction mix per kernel (cuobjdump -sass, sm_100a)\n\n" % commit)
        f.write("| kernel | instructions | top opcodes | tensor-core / TMA instructions |\n|---|---|---|---|\n")
        for short, n, ops in sorted(rows):
            special = ["%s %d" % (k, v) for k, v in ops.items() if k.startswith(("UTMA", "UTC", "TCGEN", "IMMA", "HMMA"))]
            f.write("| `%s` | %d | %s | %s |\n" % (short, n, ", ".join("%s %d" % kv for kv in ops.most_common(8)), ", ".join(special) or "none (byte gather / bit-plane work)"))
    print("wrote", len(rows), "listings to", out_dir)


if __name__ == "__main__":
    main()
