"""Map an `ncu --page source --csv` export to CUDA source lines via `nvdisasm -g` of the same cubin.
usage: linemap.py <sass-with-lineinfo.txt> <ncu_source.csv> <kernel-name-substring>"""
import csv, re, collections, sys
sass, src, kern = sys.argv[1], sys.argv[2], sys.argv[3]
lines = open(sass).read().split('\n')
start = None
for i, l in enumerate(lines):
    if l.startswith('.text.') and kern in l:
        start = i
        break
cur = None
seq = []
for l in lines[start + 1:]:
    if (l.startswith('.text.') or l.startswith('.section')) and seq:
        break
    m = re.match(r'\s*//## File "([^"]+)", line (\d+)(.*)', l)
    if m:
        cur = (m.group(1).split('/')[-1], int(m.group(2)))
        continue
    m = re.match(r'\s+/\*([0-9a-f]{4,})\*/\s+(.*?);', l)
    if m:
        seq.append((m.group(2), cur))
rows = list(csv.reader(open(src)))
hdr = rows[1]
ia = hdr.index('Instructions Executed')
isamp = hdr.index('# Samples')
data = [r for r in rows[2:] if len(r) > ia]
print(len(seq), len(data))
byline = collections.Counter(); samp = collections.Counter(); tot = 0; stot = 0
for (txt, loc), r in zip(seq, data):
    n = int(r[ia] or 0); tot += n
    sm = int(r[isamp] or 0); stot += sm
    key = loc if loc else ('?', 0)
    byline[key] += n; samp[key] += sm
print('total instr', tot, 'samples', stot)
byfile = collections.Counter(); sf = collections.Counter()
for k, n in byline.items():
    byfile[k[0]] += n; sf[k[0]] += samp[k]
for k, n in byfile.most_common():
    print(k, n, f"{n/tot*100:.1f}%", f"samples {sf[k]/stot*100:.1f}%")
for k, n in byline.most_common(int(sys.argv[4]) if len(sys.argv) > 4 else 45):
    print(k, n, f"{n/tot*100:.1f}%", f"s={samp[k]/stot*100:.1f}%")
