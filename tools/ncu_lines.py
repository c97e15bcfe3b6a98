"""Per-source-line instruction counts of one kernel from an .ncu-rep (ncu --import-source on) -- read here, no GPU needed.
  python tools/ncu_lines.py gpurun_out/prof.ncu-rep nos_b200/libnosplan.so k_plan_parallel sweep [top]
SASS addresses from the report's source page are matched to file:line through nvdisasm --print-line-info of the cubin."""
import collections, csv, os, re, subprocess, sys, tempfile


def main(rep, so, kernel, cubin_key, top=40):
    tmp = tempfile.mkdtemp()
    subprocess.run(["cuobjdump", "-xelf", "all", os.path.abspath(so)], cwd=tmp, capture_output=True)
    cub = [f for f in os.listdir(tmp) if f.startswith(cubin_key + ".")][0]
    sass = subprocess.run(["nvdisasm", "--print-line-info", os.path.join(tmp, cub)], capture_output=True, text=True).stdout.splitlines()
    start = next(i for i, l in enumerate(sass) if l.startswith(".text.") and kernel in l)
    off2line, cur = {}, None
    for l in sass[start + 1:]:
        if l.startswith(".text."):
            break
        m = re.search(r'//## File "([^"]+)", line (\d+)', l)
        if m:
            cur = (m.group(1).split("/")[-1], int(m.group(2)))
            continue
        m = re.match(r"\s*/\*([0-9a-f]{4,})\*/", l)
        if m:
            off2line[int(m.group(1), 16)] = cur
    out = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--kernel-name", "regex:" + kernel], capture_output=True, text=True).stdout
    rows = list(csv.reader(out.splitlines()))
    hdr = rows[1]
    ie, ss = hdr.index("Instructions Executed"), hdr.index("# Samples")
    base = int(rows[2][0], 16)
    agg, samp = collections.Counter(), collections.Counter()
    for r in rows[2:]:
        if len(r) <= ie or not r[ie].isdigit():
            continue
        k = off2line.get(int(r[0], 16) - base) or ("?", 0)
        agg[k] += int(r[ie]); samp[k] += int(r[ss] or 0)
    tot, stot = sum(agg.values()), sum(samp.values())
    print(f"kernel {kernel}: {tot} warp instructions, {stot} stall samples")
    srcs = {}
    for (f, ln), v in agg.most_common(top):
        if f not in srcs:
            p = os.path.join(os.path.dirname(os.path.abspath(so)), "csrc", f)
            srcs[f] = open(p).read().splitlines() if os.path.exists(p) else None
        text = srcs[f][ln - 1].strip()[:100] if srcs[f] and 0 < ln <= len(srcs[f]) else ""
        print(f"{100 * v / tot:5.1f}% inst {100 * samp[(f, ln)] / max(stot, 1):5.1f}% samples  {f}:{ln}  {text}")


if __name__ == "__main__":
    main(sys.argv[1], sys.argv[2], sys.argv[3], sys.argv[4], int(sys.argv[5]) if len(sys.argv) > 5 else 40)
