"""Attributes an `ncu --set full --import-source on` capture of one kernel to source lines / device functions.
usage: python scripts/ncu_by_line.py <report.ncu-rep> <lib.so> <kernel-mangled-substring> [top]
Joins ncu's SASS-level source page (per-instruction executed counts and stall samples) with `nvdisasm -g` of the cubin
extracted from the library (line info per instruction)."""
import collections, csv, glob, os, re, subprocess, sys, tempfile


def main():
    rep, lib, kern = sys.argv[1:4]
    top = int(sys.argv[4]) if len(sys.argv) > 4 else 40
    tmp = tempfile.mkdtemp()
    subprocess.check_call("cd %s && cuobjdump -xelf all %s > /dev/null" % (tmp, os.path.abspath(lib)), shell=True)
    cubin = glob.glob(tmp + "/*.cubin")[0]
    dis = subprocess.check_output(["nvdisasm", "-g", "-c", cubin]).decode().split("\n")
    raw = subprocess.check_output(["ncu", "-i", rep, "--page", "raw", "--csv"]).decode()
    rows = list(csv.reader(subprocess.check_output(["ncu", "-i", rep, "--page", "source", "--csv"]).decode().split("\n")))
    start = next(i for i, l in enumerate(dis) if l.startswith(".text.") and kern in l)
    end = next(i for i, l in enumerate(dis) if i > start and l.startswith("\t.section"))
    f = ln = None
    fn = "kernel"
    insts = []
    for l in dis[start:end]:
        m = re.match(r'\s*//## File "([^"]+)", line (\d+)', l)
        if m:
            f, ln = m.group(1).split("/")[-1], int(m.group(2))
            continue
        m = re.match(r"^(\S+):\s*$", l)
        if m and "_Z" in m.group(1) and not m.group(1).startswith(".text"):
            fn = m.group(1).split("$")[-1]
            continue
        m = re.match(r"\s+/\*([0-9a-f]{4,})\*/\s+(.*?);", l)
        if m:
            insts.append((m.group(2), f, ln, fn))
    hdr = next(r for r in rows if r and r[0] == "Address")
    data = [r for r in rows[rows.index(hdr) + 1:] if len(r) == len(hdr)]
    assert len(data) == len(insts), (len(data), len(insts))
    iex, isamp = hdr.index("Instructions Executed"), hdr.index("# Samples")
    byfn, byline, sfn, sline = (collections.Counter() for _ in range(4))
    for (txt, f, ln, fn), r in zip(insts, data):
        n, s = int(r[iex]), int(r[isamp])
        byfn[fn] += n; byline[(f, ln)] += n; sfn[fn] += s; sline[(f, ln)] += s
    tot, tots = sum(byfn.values()), sum(sfn.values())
    print("instructions executed %d, samples %d, SASS instructions %d" % (tot, tots, len(insts)))
    for k, n in byfn.most_common(10):
        print("  %-44s inst %5.1f%%  samples %5.1f%%" % (k[:44], 100.0 * n / tot, 100.0 * sfn[k] / tots))
    for (f, ln), n in sline.most_common(top):
        print("  %-18s %4d  samples %5.2f%%  inst %5.2f%%" % (f, ln, 100.0 * n / tots, 100.0 * byline[(f, ln)] / tot))


if __name__ == "__main__":
    main()
