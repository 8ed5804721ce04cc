"""Counts of the SASS mnemonics that prove bulk-async copies (UBLKCP), mbarriers (SYNCS), cp.async (LDGSTS) and cluster
barriers (UCGABAR) per kernel of libmpba.so:  cuobjdump -sass mpsfm_b200/lib/libmpba.so | python scripts/sass_grep.py"""
import collections
import re
import sys

want = ("UBLKCP", "SYNCS", "LDGSTS", "UCGABAR", "UTMALDG", "UTMASTG", "DFMA", "DMUL", "DADD")
fn, counts = None, collections.defaultdict(collections.Counter)
for line in sys.stdin:
    m = re.search(r"Function : (\S+)", line)
    if m:
        name = m.group(1)
        k = re.search(r"\d+(k_[a-z_0-9]+?)(E|I[LE])", name)
        fn = k.group(1) if k else name[:40]
        t = re.search(r"ILi(\d+)E", name)
        if t:
            fn += f"<{t.group(1)}>"
        continue
    if fn is None:
        continue
    for w in want:
        if re.search(r"\b" + w + r"[\w.]*", line):
            counts[fn][w] += 1
print("kernel".ljust(28) + "".join(w.rjust(9) for w in want))
for fn in sorted(counts):
    print(fn.ljust(28) + "".join(str(counts[fn][w]).rjust(9) for w in want))
