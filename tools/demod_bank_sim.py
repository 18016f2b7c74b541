"""Shared-memory bank conflicts of demod.cu's deinterleave scatter (STS wavefronts per store instruction of one warp = four
8-lane groups), from the real tables of w80211_fill_tables, for a given staging-row stride and optional per-group XOR of
the row index; brute force over even strides and XORs for the 256-QAM table.  CPU only."""
import ctypes as C, numpy as np, sys, itertools
sys.path.insert(0, __import__('os').path.dirname(__import__('os').path.dirname(__import__('os').path.abspath(__file__))))
import w80211rx
host=w80211rx.host
# the leading fields of W80211Tables (csrc/common.cuh); the tables after `map` are not needed here
TAB_DTYPE = np.dtype([("pilot_p", "<f4", 128), ("ltf_nl", "<f4", 64), ("sig_demap", "i1", 64), ("bin_j_l", "i1", 64),
                      ("bin_j_nl", "i1", 64), ("sigb_deint", "u1", 52), ("sigb_int", "u1", 52),
                      ("map", "<u2", (19, 416))], align=True)
L=host.load(); raw=np.zeros(1<<20,np.uint8); L.w80211_fill_tables.argtypes=[C.c_void_p]; L.w80211_fill_tables(raw.ctypes.data)
t=raw[:TAB_DTYPE.itemsize].view(TAB_DTYPE)[0]
cases={"vht256":(8,8,False),"ht64":(7,6,False),"ht16":(6,4,False),"htqpsk":(5,2,False),"l64":(3,6,True),"l16":(2,4,True)}
def sim(tabid, nb, legacy, base, xor):
    m=t["map"][tabid].astype(int); binj=(t["bin_j_l"] if legacy else t["bin_j_nl"]).astype(int)
    tot=0; n=0
    for k2 in range(8):
        for b in range(nb):
            banks={}
            for grp in range(4):
                for g in range(8):
                    j=binj[g+8*k2]
                    if j<0: continue  # pad slot writes: ignore (same address -> broadcast-ish)
                    addr=base[grp]+(int(m[j*nb+b])^xor[grp])
                    banks.setdefault(addr%32,set()).add(addr)
            if banks:
                tot+=max(len(v) for v in banks.values()); n+=1
    return tot/n
if __name__=="__main__":
    print("current 420 stride (nCBPS 416 + 4):", {k:round(sim(v[0],v[1],v[2],[g*420 for g in range(4)],[0]*4),2) for k,v in cases.items()})
    print("old 416 stride:", {k:round(sim(v[0],v[1],v[2],[g*416 for g in range(4)],[0]*4),2) for k,v in cases.items()})
    if "--search" not in sys.argv:
        sys.exit(0)
    best=[]
    for stride in range(416,450,2):
        for xs in itertools.product(range(0,32,2),repeat=3):
            xor=[0]+list(xs)
            base=[g*stride for g in range(4)]
            s=sim(8,8,False,base,xor)
            best.append((s,stride,xor))
        best.sort(key=lambda x:x[0]); best=best[:5]
    print(best)
