"""Opcode histograms of the decoder kernel's row bodies from the SASS of the shipped library:
    python tools/sass_row_bodies.py > profiles/r02_sass_row_bodies.txt
A row body is the code from one class label of the step dispatch to the branch back to the loop head; it is recognised by its
mbarrier wait (SYNCS.PHASECHK) and arrive (SYNCS.ARRIVE). Needs cuobjdump (CUDA toolkit), no GPU."""
import collections
import os
import re
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
LIB = os.path.join(ROOT, "srsran_project_23.5_b200", "lib", "libsrsran_ldpc_cuda.so")
KERNEL = "ldpc_decode_tmem_kernelILi1ELi1ELb0ELi1EE"  # <BG1, generic 0.8, decode-only, Z = 384>


def main():
    sass = subprocess.run(["cuobjdump", "-sass", LIB], capture_output=True, text=True).stdout
    ins, on = [], False
    for line in sass.splitlines():
        if "Function :" in line:
            on = KERNEL in line
        m = re.match(r"\s*/\*[0-9a-f]{4,6}\*/\s+(@!?U?P\d+\s+)?([A-Z0-9_]+)(\S*)\s*(.*?);", line)
        if on and m:
            ins.append((m.group(2), m.group(2) + m.group(3), m.group(4), m.group(1) is not None))
    total = collections.Counter(i[0] for i in ins)

    def fmt(c, n=18):
        return ", ".join("%s %d" % kv for kv in c.most_common(n))

    print("SASS opcode histograms, ldpc_decode_tmem_kernel<BG1, generic 0.8, decode-only, Z = 384> (cuobjdump -sass of the shipped")
    print("library, end of round 2; tools/sass_row_bodies.py). Whole kernel: %d instructions. LDTM / STTM = tcgen05.ld / tcgen05.st" % len(ins))
    print("(messages in tensor memory); SYNCS = mbarrier arrive / try_wait (the split row barrier); LDCU / U* = uniform datapath")
    print("(step descriptors, edge records, TMEM addresses); no UTC*MMA (not a contraction); HMNMX2 / VIMNMX3 = min-sum tree;")
    print("HFMA2 / HADD2 = scaled-domain message arithmetic.\n")
    print("whole kernel: " + fmt(total, 30) + "\n")
    # bodies: from the instruction after a branch to the unconditional branch that follows an arrive
    start, seen_wait, seen_arrive = 0, False, False
    for i, (op, full, _, predicated) in enumerate(ins):
        if full.startswith("SYNCS.PHASECHK"):
            seen_wait = True
        if full.startswith("SYNCS.ARRIVE"):
            seen_arrive = True
        if full in ("BRA", "EXIT", "RET") and not predicated:
            if seen_wait and seen_arrive:
                body = ins[start:i + 1]
                # cut the dispatch chain in front (compare / branch pairs) off: the body starts behind the last BRA.U
                k = max([j for j, b in enumerate(body) if b[1] == "BRA.U"] + [-1]) + 1
                body = body[k:]
                c = collections.Counter(b[0] for b in body)
                if c["LDG"]:
                    continue  # the last body's segment runs on into the CRC sweep behind the loop
                print("row body with %d LDS (values entering the check rows), %d instructions: %s" % (c["LDS"], len(body), fmt(c)))
            start, seen_wait, seen_arrive = i + 1, False, False


if __name__ == "__main__":
    sys.exit(main())
