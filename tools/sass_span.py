"""Address span / instruction count of the SASS attributed (outermost inlining site) to a range of
source lines of one file, from an `nvdisasm -g -c` listing.
usage: sass_span.py <listing> <file substring> <first line> <last line>"""
import re
import sys

lst, fsub, l0, l1 = sys.argv[1], sys.argv[2], int(sys.argv[3]), int(sys.argv[4])
cur_outer = None
addrs = []
for l in open(lst):
    m = re.match(r'\s*//## File "([^"]+)", line (\d+)(?: inlined at "([^"]+)", line (\d+))?', l)
    if m:
        # nvdisasm prints the innermost location first, then "inlined at" chains on following lines
        f, ln = m.group(1), int(m.group(2))
        if m.group(3):
            f2, ln2 = m.group(3), int(m.group(4))
        cur_outer = (f, ln)
        if m.group(3):
            cur_outer = (f2, ln2)
        continue
    m2 = re.match(r'\s*//## +inlined at "([^"]+)", line (\d+)', l)
    if m2:
        cur_outer = (m2.group(1), int(m2.group(2)))
        continue
    m3 = re.match(r"\s*/\*([0-9a-f]{4,6})\*/\s+\S", l)
    if m3 and cur_outer and fsub in cur_outer[0] and l0 <= cur_outer[1] <= l1:
        addrs.append(int(m3.group(1), 16))
if addrs:
    print(f"{len(addrs)} instructions, addresses {min(addrs):#x}..{max(addrs):#x} span {(max(addrs) - min(addrs)) / 1024:.1f} KB, "
          f"code bytes {len(addrs) * 16 / 1024:.1f} KB")
else:
    print("no instructions")
