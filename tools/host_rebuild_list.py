import numpy as np, sys
import os; sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from mv3d_b200.sweeps import hdl64_sweep
s=hdl64_sweep(0)
x,y,z=s[:,0],s[:,1],s[:,2]
X=np.floor((x-0.0)/0.1).astype(np.int64); Y=np.floor((y+40.0)/0.1).astype(np.int64); Z=np.floor((z+4.2)/0.2).astype(np.int64)
keep=(X>=1)&(X<=704)&(Y>=1)&(Y<=800)&(Z>=0)&(Z<=25)
cell=((X-1)*800+(Y-1))[keep]
T=340; tiles=(704*800+T-1)//T
tile=cell//T
cnt=np.bincount(tile,minlength=tiles)
toff=np.zeros(tiles+1,np.uint32); toff[1:]=np.cumsum(cnt)
n=int(toff[-1])
idx=np.full(n,0xFFFFFFFF,np.uint32)
uc=np.unique(cell)
ut=uc//T
# distinct cells of tile t take the first slots of the tile's range
order=np.argsort(ut,kind='stable'); uc=uc[order]; ut=ut[order]
start=toff[ut]; rank=np.arange(len(uc))-np.searchsorted(ut,ut,side='left')
idx[start+rank]=uc
print('kept points',n,'distinct cells',len(uc),'tiles',tiles)
with open(sys.argv[1] if len(sys.argv) > 1 else '/tmp/list.bin', 'wb') as f:
    np.array([n,tiles,T],np.uint32).tofile(f); toff.tofile(f); idx.tofile(f)
