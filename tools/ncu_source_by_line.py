"""Joins the SASS view of an `ncu --page source --csv` export (stall samples per instruction) with `nvdisasm -g` line info of the
same cubin: stall samples per CUDA source line of one kernel.
    python tools/ncu_source_by_line.py <source.csv.gz> <nvdisasm -g output> <mangled kernel name> <demangled substring> [top n]"""
import csv,gzip,re,collections,sys
src_csv, sass_file, mangled, kname_sub = sys.argv[1:5]
topn = int(sys.argv[5]) if len(sys.argv)>5 else 40
rows=list(csv.reader(gzip.open(src_csv,'rt')))
kern=[];cur=None
for r in rows:
    if r and r[0]=='Kernel Name':
        cur={'name':r[1],'rows':[]}; kern.append(cur)
    elif r and r[0].startswith('0x'):
        cur['rows'].append(r)
# pick the kernel instance with most samples matching name
cands=[k for k in kern if kname_sub in k['name']]
k=max(cands,key=lambda k:sum(int(r[4]) for r in k['rows']))
print(k['name'], 'samples', sum(int(r[4]) for r in k['rows']), 'sass', len(k['rows']))
lines=open(sass_file).read().split('\n')
start=next(i for i,l in enumerate(lines) if l.startswith('.text.'+mangled+':'))
end=next(i for i in range(start+1,len(lines)) if lines[i].startswith('.text.') or lines[i].startswith('.section') )
seq=[];cur=None
for l in lines[start:end]:
    m=re.search(r'//## File "([^"]+)", line (\d+)(.*)',l)
    if m:
        cur=(m.group(1).split('/')[-1],int(m.group(2))); continue
    m=re.match(r'\s+/\*[0-9a-f]+\*/\s+(.*?);',l)
    if m:
        seq.append((cur,m.group(1).strip()))
print('nvdisasm instr',len(seq))
n=min(len(seq),len(k['rows']))
mism=0
by=collections.Counter(); inst=collections.Counter()
for i in range(n):
    op1=[o for o in k['rows'][i][1].split() if not o.startswith('@')][0]
    op2=[o for o in seq[i][1].split() if not o.startswith('@')][0]
    if op1.split('.')[0]!=op2.split('.')[0]: mism+=1
    by[seq[i][0]]+=int(k['rows'][i][4]); inst[seq[i][0]]+=int(k['rows'][i][5])
print('opcode mismatches',mism)
tot=sum(by.values())
for key,c in by.most_common(topn):
    print(f'{key[0]:24s} {key[1]:5d}  {100*c/tot:5.1f}%  inst {inst[key]}')
