import sys,re,subprocess,csv,os,tempfile
obj,rep,kname=sys.argv[1],sys.argv[2],sys.argv[3]
d=tempfile.mkdtemp(); 
subprocess.run(['cuobjdump','-xelf','all',os.path.abspath(obj)],cwd=d,capture_output=True)
cub=[f for f in os.listdir(d) if f.endswith('.cubin')][0]
dis=subprocess.run(['nvdisasm','-c',os.path.join(d,cub)],capture_output=True,text=True).stdout
# find kernel section
sec=None; funcs=[]  # (offset,name)
cur=None
for line in dis.splitlines():
    m=re.match(r'\s*\.text\.(\S+):',line)
    if m: sec=m.group(1); continue
    if sec is None or kname not in sec: continue
    m=re.match(r'^(\S+):\s*$',line)
    if m and not m.group(1).startswith('.L_'):
        cur=m.group(1); continue
    m=re.match(r'\s*/\*([0-9a-f]{4,6})\*/\s+(.*?);',line)
    if m:
        off=int(m.group(1),16)
        if cur is not None: funcs.append((off,cur)); cur=None
def short(n):
    m=re.search(r'mib(\d+)([A-Za-z_0-9]+?)E',n)
    if '$' in n:
        t=n.split('$')[-1]
        m=re.search(r'mib\d+([a-z_0-9]+?)(E|I)',t)
        return m.group(1) if m else t[:40]
    return 'KERNEL BODY'
funcs.sort()
sass=subprocess.run(['ncu','-i',rep,'--page','source','--csv','--print-source','sass'],capture_output=True,text=True).stdout
rows=list(csv.reader(sass.splitlines()))
hdr=rows[1]; data=[r for r in rows[2:] if len(r)==len(hdr)]
# a report with several launches repeats the header: keep the rows of the last launch
last=max([i for i,r in enumerate(data) if r[hdr.index('Address')]=='Address'], default=-1); data=data[last+1:]
TI=hdr.index('Thread Instructions Executed') if 'Thread Instructions Executed' in hdr else None
iE=hdr.index('Instructions Executed'); iA=hdr.index('Address'); iSm=hdr.index('# Samples')
base=int(data[0][iA],16)
import bisect
offs=[f[0] for f in funcs]
agg={}
tot=0; totS=0
for r in data:
    off=int(r[iA],16)-base; e=int(r[iE]); sm=int(r[iSm])
    i=bisect.bisect_right(offs,off)-1
    name=short(funcs[i][1]) if i>=0 else '?'
    a=agg.setdefault(name,[0,0,0,0,0]); a[0]+=1; a[1]+= (e>0); a[2]+=e; a[3]+=sm; a[4]+=(int(r[TI]) if TI is not None else 0); tot+=e; totS+=sm
print('%-28s %7s %7s %8s %8s %6s'%('function','instr','hot','exec%','stall%','lanes'))
for k,v in sorted(agg.items(),key=lambda kv:-kv[1][2]): print('%-28s %7d %7d %7.1f%% %7.1f%% %6.1f'%(k,v[0],v[1],100*v[2]/tot,100*v[3]/max(totS,1),v[4]/max(v[2],1)))
