#!/usr/bin/env python
"""Lists the memory instructions of one kernel in a cubin / .so with the scoreboard each one writes (wr), the read
barrier it holds (rd) and the scoreboards an instruction waits for (wait) -- decoded from the control bits of the
SASS encoding (bits 105-125 of every 128-bit instruction).  ptxas gives a kernel six scoreboards; loads that share one
complete together as far as a waiting instruction can tell, which is what decides how many loads are really in flight.

  python tools/sass_scoreboards.py distributed-tf-serving_b200/libctr_b200.so gather_tile_kernelILi0ELb0ELb1ELb1ELb0ELb0ELb0E [regex]
"""
import re
import subprocess
import sys


def main():
    lib, kern = sys.argv[1], sys.argv[2]
    pat = re.compile(sys.argv[3] if len(sys.argv) > 3 else r"LDG|STG|LDS|SHFL|STS|SYNCS|UBLKCP")
    out = subprocess.run(["cuobjdump", "-sass", lib], capture_output=True, text=True).stdout.split("\n")
    on, i = False, 0
    while i < len(out):
        ln = out[i]
        if "Function :" in ln:
            on = kern in ln
        m = re.match(r"\s+/\*([0-9a-f]{4,5})\*/\s+(.*?);\s+/\* (0x[0-9a-f]+) \*/", ln) if on else None
        if m and i + 1 < len(out):
            m2 = re.match(r"\s+/\* (0x[0-9a-f]+) \*/", out[i + 1])
            if m2:
                ctrl = (int(m2.group(1), 16) >> 41) & 0x1FFFFF
                wr, rd, wait = (ctrl >> 5) & 7, (ctrl >> 8) & 7, (ctrl >> 11) & 0x3F
                txt = m.group(2).strip()
                if pat.search(txt) or (len(sys.argv) > 4 and wait):
                    print(m.group(1), txt[:64].ljust(64), "wr=%s" % (wr if wr != 7 else "-"), "rd=%s" % (rd if rd != 7 else "-"),
                          "wait=%s" % ("".join(str(b) for b in range(6) if wait >> b & 1) or "-"))
                i += 2
                continue
        i += 1


if __name__ == "__main__":
    main()
