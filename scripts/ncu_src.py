"""Helpers to read `ncu --page source --csv` exports: python scripts/ncu_src.py file.csv [kernel-index]"""
import csv, sys
def load(path):
    rows=list(csv.reader(open(path)))
    secs=[]; cur=None
    for r in rows:
        if r and r[0]=="Kernel Name": cur={'name':r[1],'hdr':None,'rows':[]}; secs.append(cur); continue
        if cur is None: continue
        if cur['hdr'] is None: cur['hdr']=r; continue
        cur['rows'].append(r)
    return secs
if __name__=="__main__":
    secs=load(sys.argv[1])
    seen=set()
    for s in secs:
        if s['name'] in seen: continue
        seen.add(s['name'])
        h=s['hdr']; ia=h.index('Address'); isrc=h.index('Source'); ismp=h.index('# Samples'); iex=h.index('Instructions Executed')
        R=s['rows']; tot=sum(int(r[ismp] or 0) for r in R)
        print(s['name'][:70],'samples',tot)
        stall=[i for i,x in enumerate(h) if x.startswith('stall_') and 'Not' not in x]
        agg={h[i]:sum(int(r[i] or 0) for r in R) for i in stall}
        print(' ',sorted(agg.items(), key=lambda kv:-kv[1])[:6])
        print('  trywait sites:')
        for i,r in enumerate(R):
            if 'TRYWAIT' in r[isrc]:
                print('   ',r[ia][-5:], 'samples', int(r[ismp])+int(R[i+1][ismp]), 'iters', r[iex], 'entered', R[i-2][iex], r[isrc][40:80])
        # region sample sums between UTCHMMA first/last
        idx=[i for i,r in enumerate(R) if 'UTCHMMA' in r[isrc]]
        if idx:
            print('  samples in [first UTCHMMA-60, last UTCHMMA+40]:', sum(int(r[ismp]) for r in R[idx[0]-60:idx[-1]+40]), 'instrs', idx[-1]-idx[0]+100)
