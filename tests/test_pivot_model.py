"""Executable model of the pivot rounds that refine sharded parts (k_fb_pivot_pick / k_fb_keys_pivot /
k_fb_regroup_pivot in genometools_b200/csrc/gtsfx_b200.cu), with 3-symbol windows and a cap of 4
windows so that small texts exercise every branch: class order, the lcp of neighbours of different classes, groups cut
by specials (ordered by position), the cap.  Checked against a naive suffix comparison that follows the reference's
rules (a special is larger than every symbol, two specials compare by position, a special never matches).  CPU only;
the CUDA kernels are held to the oracle by tests/test_gpu_parity.py."""
import random
W=3; BITS=2; MAXWIN=4; POS=-1
def make_key(t,n,q):
    # W symbols from q; positions >= first special padded with ones; tag = W - avail
    avail=0
    while avail<W and q+avail<n and t[q+avail]<4: avail+=1
    x=0
    for i in range(W):
        s = t[q+i] if i<avail else 3
        x=(x<<BITS)|s
    return (x<<8)|(W-avail)
def key_common(ka,kb):
    x=ka^kb
    # count leading equal symbols in the 56-bit... here W*BITS bits above the tag
    common=W
    for i in range(W):
        sh=8+BITS*(W-1-i)
        if ((ka>>sh)&3)!=((kb>>sh)&3): common=i;break
    aa=W-(ka&0xff); ab=W-(kb&0xff)
    return min(common,aa,ab)
def naive_cmp_lcp(t,n,a,b):
    l=0
    while True:
        sa = t[a+l] if a+l<n else 255
        sb = t[b+l] if b+l<n else 255
        if sa>=4 or sb>=4:
            if sa>=4 and sb>=4: return (-1 if a<b else 1), l
            return (1 if sa>=4 else -1), l
        if sa!=sb: return (-1 if sa<sb else 1), l
        l+=1
def run(t,pl,rng,pivot_first_round=False):
    n=len(t)
    # buckets
    from collections import defaultdict
    b=defaultdict(list)
    for p in range(n):
        if all(p+i<n and t[p+i]<4 for i in range(pl)):
            b[tuple(t[p:p+pl])].append(p)
    total_rounds=0
    results=[]  # (sorted list, lcp list within bucket)
    for code in sorted(b):
        mem=b[code][:]
        rng.shuffle(mem)
        m=len(mem)
        sa=mem[:]; lcp=[None]*m
        groups=[(0,m,pl)] if m>1 else []
        rounds=0
        while groups:
            rounds+=1
            newgroups=[]
            for (s,eend,d) in groups:
                ms=sa[s:eend]
                if d==POS:
                    ms.sort(); sa[s:eend]=ms; continue
                q=rng.choice(ms)
                recs=[]
                for p in ms:
                    e=0
                    while True:
                        ka=make_key(t,n,p+d+e*W); kb=make_key(t,n,q+d+e*W)
                        if ka!=kb or (ka&0xff)!=0 or e==MAXWIN-1: break
                        e+=1
                    c = e if ka<kb else (MAXWIN if ka==kb else 2*MAXWIN-e)
                    recs.append((c,ka,p,e,kb))
                recs.sort(key=lambda r:(r[0],r[1]))   # stable
                heads=[]
                for j,(c,k,p,e,kp) in enumerate(recs):
                    sa[s+j]=p
                    hd=True
                    if j>0:
                        c0,k0,p0,e0,kp0=recs[j-1]
                        if c0==c:
                            D=d+e*W
                            if k0!=k: lcp[s+j]=D+key_common(k0,k)
                            elif (k&0xff)!=0: lcp[s+j]=D+(W-(k&0xff)); hd=False
                            else: hd=False
                        else:
                            if e0<e: S=recs[j-1]
                            elif e<e0: S=recs[j]
                            else: S=recs[j] if c0==MAXWIN else recs[j-1]
                            lcp[s+j]=d+S[3]*W+key_common(S[1],S[4])
                    heads.append(hd)
                # subgroups
                j=0
                while j<len(recs):
                    k2=j+1
                    while k2<len(recs) and not heads[k2]: k2+=1
                    if k2-j>1:
                        c,k,p,e,kp=recs[j]
                        nd = POS if (k&0xff) else d+(e+1)*W
                        newgroups.append((s+j,s+k2,nd))
                    j=k2
            groups=newgroups
            assert rounds<10000
        total_rounds=max(total_rounds,rounds)
        # verify
        for j in range(1,m):
            cmp,l=naive_cmp_lcp(t,n,sa[j-1],sa[j])
            assert cmp<0,(code,j,sa[j-1],sa[j])
            assert lcp[j]==l,(code,j,sa[j-1],sa[j],lcp[j],l)
    return total_rounds


def test_pivot_round_model_matches_naive_order_and_lcp():
    rng = random.Random(1)
    worst = 0
    for trial in range(900):
        kind = trial % 6
        n = rng.randint(1, 120)
        if kind == 0:
            t = [rng.randint(0, 3) for _ in range(n)]
        elif kind == 1 or kind == 5:
            t = [rng.randint(0, 1) for _ in range(n)]
        elif kind == 2:
            u = [rng.randint(0, 3) for _ in range(rng.randint(1, 7))]
            t = (u * 60)[:n]
        elif kind == 3:
            t = [3] * n
        else:
            u = [rng.randint(0, 2) for _ in range(rng.randint(5, 30))]
            t = (u + [254] + u + [rng.randint(0, 3) for _ in range(5)] + u)[:max(n, 10)]
        if not (kind == 3 and trial % 12 == 3):              # keep some pure homopolymers
            for _ in range(rng.randint(0, 4)):
                t[rng.randrange(len(t))] = rng.choice([254, 255])
        worst = max(worst, run(t, rng.randint(1, 3), rng))
    assert worst <= 40          # 120-symbol runs: pivot rounds stay logarithmic (a window per round would need 40)
