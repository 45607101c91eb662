"""Timing of cabac_decode_slices_dev (decode_slice_kernel) on copies of the fixture slices: dev tool, B200 only."""
import sys, time, numpy as np, torch
sys.path.insert(0,'tests'); sys.path.insert(0,'tests/golden'); sys.path.insert(0,'.')
import entropy_coding_b200 as E
from entropy_coding_b200.types import *
import test_syntax as T
eng=E.CabacEngine(0)
stream=torch.cuda.Stream(device=0); torch.cuda.set_stream(stream); eng.set_stream(stream.cuda_stream)
items=T._substreams()
for n in (64, 1024, 8192, 65536):
    order=[k%len(items) for k in range(n)]
    params,sps,ctx_in=np.zeros(n,PARAMS_DT),np.zeros(n,SLICE_DT),np.zeros((n,379),CTX_DT)
    data,off=[],[0]; bins=0
    for j,k in enumerate(order):
        tag,s,sp=items[k]
        params[j]=(s['qp'],s['init_id'],s['flags']|SUB_TERMINATE|SUB_CTX_IN,0); sps[j]=sp
        ctx_in[j]['state0'],ctx_in[j]['state1']=s['ctx_in'][:,0],s['ctx_in'][:,1]
        data.append(s['payload']); off.append(off[-1]+s['payload'].size); bins+=s['num_bins']
    cap=E.CabacEngine.slice_capacities(sps)
    dev=lambda a: torch.from_numpy(np.ascontiguousarray(a).view(np.uint8).reshape(-1)).cuda()
    d_data=dev(np.concatenate(data)); d_off=dev(np.asarray(off,np.uint64)); d_p=dev(params); d_sp=dev(sps); d_ci=dev(ctx_in)
    d_cu=dev(cap['cu']); d_tu=dev(cap['tu']); d_ctu=dev(cap['ctu']); d_co=dev(cap['coeff'])
    cus=torch.zeros(int(cap['cu'][-1])*32,dtype=torch.uint8,device='cuda'); tus=torch.zeros(int(cap['tu'][-1])*16,dtype=torch.uint8,device='cuda')
    tuf=torch.zeros(int(cap['tu'][-1]),dtype=torch.uint8,device='cuda'); ctus=torch.zeros(int(cap['ctu'][-1])*32,dtype=torch.uint8,device='cuda')
    co=torch.zeros(int(cap['coeff'][-1]),dtype=torch.int32,device='cuda'); cnt=torch.zeros(n*3,dtype=torch.int32,device='cuda')
    nb=torch.zeros(n,dtype=torch.int32,device='cuda'); st=torch.zeros(n,dtype=torch.int32,device='cuda')
    def run(): eng.decode_slices_dev(d_data,d_off,d_p,d_sp,n,d_cu,d_tu,d_ctu,d_co,cus,tus,tuf,ctus,co,counts=cnt,num_bins=nb,ctx_in=d_ci,status=st)
    for _ in range(3): run()
    torch.cuda.synchronize()
    e0,e1=torch.cuda.Event(enable_timing=True),torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(5): run()
    e1.record(); torch.cuda.synchronize()
    ms=e0.elapsed_time(e1)/5
    assert int(st.abs().sum())==0 and int(nb.sum())==bins
    print(f"decode_slices n={n}: {ms:.3f} ms, {bins/ms/1e6:.3f} Gbins/s, {off[-1]/ms/1e6:.3f} GB/s in, coeff out {int(cap['coeff'][-1])*4/ms/1e6:.1f} GB/s cap", flush=True)

# Reference point: the SAME blocks as residual-only substreams through the lockstep residual decoder
# (cabac_decode_residual_dev, TU list as input) - what the per-block work costs without the CTU-level parser.
from entropy_coding_b200.types import TU_DT, SUB_ALIGN
for n in (8192,):
    order=[k%len(items) for k in range(n)]
    tus=[];tuoff=[0];par=np.zeros(n,PARAMS_DT);coeffs=[];cbase=0
    for j,k in enumerate(order):
        tag,s,sp=items[k]
        t=s['tus'].copy(); t['coeff_offset']+=cbase; t['flags']&=~np.uint8(0)  # flags as recorded (TS decided)
        tus.append(t); tuoff.append(tuoff[-1]+len(t)); coeffs.append(s['coeffs']); cbase+=s['coeffs'].size
        par[j]=(s['qp'],s['init_id'],(s['flags']&7)|SUB_TERMINATE|SUB_ALIGN,0)
    tus=np.concatenate(tus); coeffs=np.concatenate(coeffs).astype(np.int32); tuoff=np.asarray(tuoff,np.uint64)
    cap=np.concatenate([[0],np.cumsum([max(64,int(items[k][1]['payload'].size*2+64)) for k in order])]).astype(np.uint64)
    r=eng.encode_residual(tus,tuoff,par,coeffs,cap)
    sizes=r['sizes'].astype(np.int64); ent=np.concatenate([[0],np.cumsum(sizes)]).astype(np.uint64)
    packed=np.concatenate([r['out'][int(cap[i]):int(cap[i])+int(sizes[i])] for i in range(n)])
    bins=int(r['num_bins'].astype(np.int64).sum())
    dev=lambda a: torch.from_numpy(np.ascontiguousarray(a).view(np.uint8).reshape(-1)).cuda()
    d_in,d_ent,d_tus,d_tuoff,d_par=dev(packed),dev(ent),dev(tus),dev(tuoff),dev(par)
    d_co=torch.zeros(coeffs.size,dtype=torch.int32,device='cuda'); d_nb=torch.zeros(n,dtype=torch.int32,device='cuda'); d_st=torch.zeros(n,dtype=torch.int32,device='cuda')
    def run2(): eng.decode_residual_dev(d_in,d_ent,d_tus,d_tuoff,d_par,n,d_co,d_nb,status=d_st)
    for _ in range(3): run2()
    torch.cuda.synchronize()
    e0,e1=torch.cuda.Event(enable_timing=True),torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(5): run2()
    e1.record(); torch.cuda.synchronize()
    ms=e0.elapsed_time(e1)/5
    assert int(d_st.abs().sum())==0 and bool((d_co.cpu().numpy()==coeffs).all())
    print(f"decode_residual (same blocks, residual-only substreams) n={n}: {ms:.3f} ms, {bins/ms/1e6:.3f} Gbins/s ({bins} residual bins)", flush=True)
