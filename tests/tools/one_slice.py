import sys, os, numpy as np
sys.path.insert(0,'tests'); sys.path.insert(0,'tests/golden'); sys.path.insert(0,'.')
import entropy_coding_b200 as E
import test_syntax as T
from entropy_coding_b200.types import *
idx=int(sys.argv[1]) if len(sys.argv)>1 else 0
tag,s,sp=T._substreams()[idx]
eng=E.CabacEngine(0)
params=np.array([(s['qp'],s['init_id'],s['flags']|SUB_TERMINATE|SUB_CTX_IN,0)],PARAMS_DT)
ctx_in=np.zeros((1,379),CTX_DT); ctx_in[0]['state0'],ctx_in[0]['state1']=s['ctx_in'][:,0],s['ctx_in'][:,1]
r=eng.decode_slices(s['payload'],np.array([0,s['payload'].size],np.uint64),params,np.array([sp],SLICE_DT),ctx_in=ctx_in,raise_on_status=False)
print(tag,'status',r['status'],'bins',r['num_bins'],'expected',s['num_bins'],'counts',r['counts'])
