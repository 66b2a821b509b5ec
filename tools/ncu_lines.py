"""Development helper: executed warp instructions per CUDA source line of the first kernel in an .ncu-rep
(argv[1]; needs -lineinfo and --import-source on); argv[2] = number of lines to print."""
import collections
import csv
import subprocess
import sys

out = subprocess.run(["ncu", "-i", sys.argv[1], "--page", "source", "--csv", "--print-source", "cuda,sass"],
                     capture_output=True, text=True).stdout
cur, agg, src, ops = None, collections.Counter(), {}, collections.Counter()
iI = None
for r in csv.reader(out.splitlines()):
    if not r:
        continue
    if r[0] == "File Path":
        cur = r[1].split("/")[-1]
    elif r[0] == "Line No":
        iI = r.index("Instructions Executed")
    elif r[0].isdigit():
        try:
            agg[(cur, int(r[0]))] += int(r[iI])
        except ValueError:
            pass
        src[(cur, int(r[0]))] = r[1]
    elif r[0] == "" and len(r) > 3 and iI is not None:
        try:
            ops[r[3].split()[0] if not r[3].strip().startswith("@") else r[3].split()[1]] += int(r[iI])
        except (ValueError, IndexError):
            pass
tot = sum(agg.values())
print("total warp instructions", tot)
for k, v in sorted(agg.items(), key=lambda kv: -kv[1])[:int(sys.argv[2]) if len(sys.argv) > 2 else 40]:
    print("%-22s %4d %5.1f%%  %s" % (k[0][:22], k[1], 100 * v / tot, src[k][:100]))
print("opcodes:")
for k, v in sorted(ops.items(), key=lambda kv: -kv[1])[:30]:
    print("  %-24s %5.1f%%" % (k, 100 * v / tot))
