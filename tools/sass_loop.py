"""Instruction histogram of the hot loop of a kernel from `cuobjdump -sass` output: the innermost backward branch
whose body holds more than 90 DFMA. Usage: cuobjdump -sass build/mlsa_seg.o > x.sass; python tools/sass_loop.py x.sass k_mlsa_v3ILi35 [regex of instructions to list]"""
import re,sys
from collections import Counter
lines=open(sys.argv[1]).read().split('\n')
pat=sys.argv[2]
cur=None; ins=[]
for l in lines:
    if 'Function :' in l:
        cur=l.split(':')[1].strip(); continue
    if cur and pat in cur:
        m=re.match(r'\s+/\*([0-9a-f]{4,})\*/\s+(.*?);',l)
        if m: ins.append((int(m.group(1),16),m.group(2).strip()))
loops=[]
for a,t in ins:
    mm=re.search(r'BRA\S*\s+(?:\S+,\s*)?(0x[0-9a-f]+)',t)
    if mm and int(mm.group(1),16)<a: loops.append((int(mm.group(1),16),a))
# the sample loop: the loop with the most DFMA density
best=None
for lo,hi in loops:
    body=[t for a,t in ins if lo<=a<=hi]
    nd=sum(1 for t in body if 'DFMA' in t)
    if nd>90 and (best is None or hi-lo<best[1]-best[0]): best=(lo,hi)
lo,hi=best
body=[(a,t) for a,t in ins if lo<=a<=hi]
c=Counter(re.sub(r'^@!?U?P\d+\s+','',t).split()[0].split('.')[0] for a,t in body)
print(hex(lo),hex(hi),len(body),dict(c.most_common(30)))
if len(sys.argv)>3:
    for a,t in body:
        if re.search(sys.argv[3],t): print(hex(a),t)
