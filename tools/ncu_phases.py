"""Warp instructions and shared-memory wavefronts of one kernel by line of its own source, from an
ncu capture (--set full --import-source on) joined with nvdisasm's inline line info of the object file.

    python tools/ncu_phases.py <file.ncu-rep> <object.o> <mangled-name-substring> <units> [line:label ...]

`units` divides the counts (tiles or CTAs of the launch). The ncu source page lists SASS in address order
and the compiler moves code across source lines, so sums between barriers by address mislead; here every
instruction is booked on the outermost line of the kernel it was inlined into. `line:label` arguments
name the first line of each phase; without them every line with executed instructions is printed.
"""
import csv
import os
import re
import subprocess
import sys
import tempfile


def line_map(obj, name):
    """offset of every instruction of the function -> (kernel source line, innermost file:line)"""
    tmp = tempfile.mkdtemp(prefix="ncu_phases_")
    subprocess.run(["cuobjdump", "-xelf", "all", os.path.abspath(obj)], cwd=tmp, check=True, capture_output=True)
    cubin = [f for f in os.listdir(tmp) if f.endswith(".cubin")][0]
    txt = subprocess.run(["nvdisasm", "-gi", "-c", os.path.join(tmp, cubin)], capture_output=True, text=True).stdout
    out, inside, outer, inner, fresh = {}, False, None, None, True
    for ln in txt.splitlines():
        if ln.startswith(".text."):
            inside = name in ln
            continue
        if not inside:
            continue
        m = re.match(r'\s*//## File "([^"]+)", line (\d+)(.*)', ln)
        if m:
            if fresh:
                inner = (os.path.basename(m.group(1)), int(m.group(2)))
                fresh = False
            if "inlined at" not in m.group(3):
                outer = int(m.group(2))
            continue
        m = re.match(r"\s*/\*([0-9a-f]{4,})\*/", ln)
        if m:
            out[int(m.group(1), 16)] = (outer, inner)
            fresh = True
    return out


def main():
    rep, obj, name, units = sys.argv[1], sys.argv[2], sys.argv[3], float(sys.argv[4])
    marks = sorted((int(a.split(":")[0]), a.split(":", 1)[1]) for a in sys.argv[5:])
    lm = line_map(obj, name)
    raw = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--print-source", "sass"],
                         capture_output=True, text=True).stdout
    rows = list(csv.reader(raw.splitlines()))
    hdr = next(r for r in rows if "Instructions Executed" in r)
    ia, ii, iw, ic = hdr.index("Address"), hdr.index("Instructions Executed"), hdr.index("L1 Wavefronts Shared"), \
        hdr.index("L1 Wavefronts Shared Excessive")
    body = [r for r in rows[rows.index(hdr) + 1:] if len(r) > ii and r[ia].startswith("0x")]
    base = int(body[0][ia], 16)
    by_line = {}
    for r in body:
        outer, _ = lm.get(int(r[ia], 16) - base, (None, None))
        a = by_line.setdefault(outer, [0, 0, 0, 0])
        a[0] += int(r[ii] or 0);  a[1] += int(r[iw] or 0);  a[2] += int(r[ic] or 0);  a[3] += 1
    tot = [sum(v[k] for v in by_line.values()) for k in range(3)]
    print(f"total per unit: {tot[0] / units:.0f} warp instructions, {tot[1] / units:.0f} shared wavefronts ({tot[2] / units:.0f} excessive)")
    if marks:
        acc = {}
        for line, v in by_line.items():
            label = "(no line)"
            for first, lab in marks:
                if line is not None and line >= first:
                    label = lab
            a = acc.setdefault(label, [0, 0, 0, 0])
            for k in range(4):
                a[k] += v[k]
        for lab in [m[1] for m in marks] + ["(no line)"]:
            if lab in acc:
                v = acc[lab]
                print(f"{lab:28s} {v[0] / units:8.0f} instr {100 * v[0] / tot[0]:5.1f} %  {v[1] / units:7.0f} wavefronts "
                      f"({v[2] / units:.0f} excessive)  {v[3]} SASS lines")
    else:
        for line in sorted(k for k in by_line if k is not None):
            v = by_line[line]
            if v[0]:
                print(f"line {line:4d} {v[0] / units:8.1f} instr  {v[1] / units:7.1f} wavefronts ({v[2] / units:.1f} excessive)  {v[3]} SASS lines")


if __name__ == "__main__":
    main()
