import re,collections,sys
d=collections.defaultdict(lambda: {"base":[], "new":[]})
cur=None
for l in open(sys.argv[1]):
    if l.startswith("---"): cur=l.split()[1]; continue
    m=re.match(r"(\S+)\s+b256 ms=([\d.]+)",l)
    if m: d[m.group(1)][cur].append(float(m.group(2)))
for k,v in d.items():
    b=min(v["base"]); n=min(v["new"]); print(f"{k:24s} base {b*1000:7.1f} us  new {n*1000:7.1f} us  {100*(n/b-1):+5.1f} %")
