import csv,re,collections,sys
sass,csvf,prefer=sys.argv[1],sys.argv[2],sys.argv[3]
lines=open(sass).read().split('\n')
cur=('?',0)
instrs=[]
for l in lines:
    if '//## File' in l:
        locs=re.findall(r'"([^"]+)", line (\d+)',l)
        pick=None
        for f,n in locs:
            if f.endswith(prefer): pick=(prefer,int(n)); break
        if pick is None and locs: pick=(locs[-1][0].split('/')[-1],int(locs[-1][1]))
        cur=pick
        continue
    m=re.match(r'\s*/\*([0-9a-f]{4,5})\*/\s+(.*?);',l)
    if m: instrs.append((cur[0],cur[1],m.group(2)))
rows=list(csv.reader(open(csvf)))
hdr=rows[1]; ix={n:i for i,n in enumerate(hdr)}; data=rows[2:]
print(len(instrs),len(data))
agg=collections.defaultdict(lambda:[0,0,0,collections.Counter()])
stall_cols=[n for n in hdr if n.startswith('stall_') and 'Not Issued' not in n]
tot_s=tot_i=0
for (f,ln,txt),r in zip(instrs,data):
    s=int(r[ix['# Samples']]); ie=int(r[ix['Instructions Executed']])
    a=agg[(f,ln)]; a[0]+=s;a[1]+=ie;a[2]+=1
    for c in stall_cols:
        v=int(r[ix[c]] or 0)
        if v:a[3][c]+=v
    tot_s+=s;tot_i+=ie
print('total samples',tot_s,'warp instr',tot_i)
src={}
import glob,os
for fn in glob.glob('/root/repo/rdf_fusion_b200/csrc/*'):
    try: src[os.path.basename(fn)]=open(fn).read().split('\n')
    except: pass
N=int(sys.argv[4]) if len(sys.argv)>4 else 45
for (f,ln),a in sorted(agg.items(),key=lambda kv:-kv[1][0])[:N]:
    st=', '.join('%s %d'%(k[6:],v) for k,v in a[3].most_common(3))
    code=src[f][ln-1].strip()[:100] if f in src and ln-1<len(src[f]) else ''
    print('%5.1f%%s %5.1f%%i %s:%d [%s] %s'%(100*a[0]/tot_s,100*a[1]/tot_i,f,ln,st,code))
