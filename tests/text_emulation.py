"""Lane-by-lane Python restatement of deep-ctr_b200/csrc/text.cu (same 4 KB index tiles, 32-byte chunks, ballots and
prefix counts) -- test infrastructure: lets the CPU suite check the TOKENISATION RULES of the device parser against the
reference parsers (oracle/text_oracle.py) without a GPU.  The CUDA code itself is checked by tests/test_widening_gpu.py."""
import numpy as np
NL=10
def is_ws(c): return c==32 or 9<=c<=13
def index_lines(t, max_lines):
    nb=len(t); tiles=(nb+4095)//4096
    def mask(p):
        if p>=nb: return 0
        b=[t[p+k] if p+k<nb else NL for k in range(16)]
        prev=t[p-1] if p>0 else NL
        m=0
        for k in range(16):
            if b[k]!=NL and prev==NL and p+k<nb: m|=1<<k
            prev=b[k]
        return m
    counts=[sum(bin(mask(tb*4096+th*16)).count("1") for th in range(256)) for tb in range(tiles)]
    off=np.concatenate([[0],np.cumsum(counts)]).astype(int)
    total=int(off[-1]); ls=np.zeros(max_lines+1,dtype=np.int64)
    if total<=max_lines: ls[total]=nb
    for tb in range(tiles):
        idx=off[tb]
        for th in range(256):
            p=tb*4096+th*16; m=mask(p)
            while m:
                k=(m&-m).bit_length()-1; m&=m-1
                if idx<=max_lines: ls[idx]=p+k
                idx+=1
    return ls,total
def parse_number(t,q,le,frac):
    get=lambda q:t[q]
    ch=get(q); neg=False
    if ch in (45,43): neg=ch==45; q+=1; ch=get(q) if q<le else NL
    v=0;n=0
    while 48<=ch<=57:
        if n<18: v=v*10+ch-48
        n+=1;q+=1; ch=get(q) if q<le else NL
    val=float(v)
    if frac and ch==46:
        q+=1; ch=get(q) if q<le else NL
        f=0;sc=1.0;nf=0
        while 48<=ch<=57:
            if nf<18: f=f*10+ch-48; sc*=10
            nf+=1;n+=1;q+=1; ch=get(q) if q<le else NL
        val+=f/sc
    return (-val if neg else val),(-v if neg else v),(-1 if n>18 else n),ch
def parse(t,ls,n_total,max_lines,criteo,skip=0,J=0,pad=-1,n_num=0,n_cat=0,offsets=None,max_vals=None):
    n=min(n_total,max_lines); F=n_num+n_cat if criteo else J
    ids=np.full((n,F),-777,dtype=np.int64); wts=np.full((n,F),-7.0,dtype=np.float32); y=np.full(n,-9.0,np.float32); status=np.zeros(n,np.int64)
    for line in range(n):
        s,le=int(ls[line]),int(ls[line+1]); count=0; bad=False; prev=9 if criteo else 32; done=False; base=s
        while base<le and not done:
            c=[t[base+l] if base+l<le else NL for l in range(32)]
            pc=[prev]+c[:31]
            if criteo:
                nl=[x==NL for x in c]
                first=nl.index(True) if any(nl) else None
                live=[(first is None or l<=first) for l in range(32)]
                tabs=[live[l] and c[l]==9 for l in range(32)]
                for l in range(32):
                    q=base+l
                    start=live[l] and q<le and pc[l]==9
                    fi=count+sum(tabs[:l])
                    if start:
                        val,iv,nd,term=parse_number(t,q,le,True)
                        if nd<=0 or term not in (9,10,13): bad=True
                        if fi==0: y[line]=val
                        elif fi<=n_num: ids[line,fi-1]=fi-1; wts[line,fi-1]=np.float32(val/max_vals[fi-1])
                        elif fi<=F:
                            g=iv+int(offsets[fi-1-n_num])
                            if g<0 or g>0x7fffffff: bad=True
                            ids[line,fi-1]=g; wts[line,fi-1]=1
                count+=sum(tabs); done=first is not None
            else:
                st=[(base+l<le) and (not is_ws(c[l])) and is_ws(pc[l]) for l in range(32)]
                for l in range(32):
                    if st[l]:
                        q=base+l; w=count+sum(st[:l])
                        val,iv,nd,term=parse_number(t,q,le,False)
                        if nd<=0 or not (is_ws(term) or (term==58 and w>0)): bad=True
                        if w==0: y[line]=iv
                        elif w>skip:
                            j=w-1-skip
                            if iv<0 or iv>0x7fffffff: bad=True
                            if j<J: ids[line,j]=iv; wts[line,j]=1
                count+=sum(st)
            prev=c[31]; base+=32
        if criteo: filled=count; stat=count+1
        else:
            filled=max(count-1-skip,0); stat=-1 if count==0 else filled
            if count==0: y[line]=0
        for j in range(filled,F): ids[line,j]=0 if criteo else pad; wts[line,j]=0
        status[line]=-2 if bad else stat
    return ids,wts,y,status
