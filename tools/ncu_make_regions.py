import re,sys
# derive regions from function markers in the sources
def lines(p): return open(p).read().split('\n')
fast=lines('/root/repo/claragenomicsanalysis_b200/csrc/poa_fast.cuh')
dev=lines('/root/repo/claragenomicsanalysis_b200/csrc/poa_device.cuh')
def find(ls,pat,start=0):
    for i in range(start,len(ls)):
        if pat in ls[i]: return i+1
    return None
R=[]
f_pred=find(fast,'__device__ __forceinline__ int32_t pred_row'); f_fm=find(fast,'struct FastMatrix') or find(fast,'__device__ __forceinline__ int32_t tilted_score_global'); f_lt=find(fast,'__device__ __forceinline__ void load_tile'); f_al=find(fast,'__device__ int32_t fast_fill_impl(')
f_pre=find(fast,'// pre-pass: row descriptors'); f_row0=find(fast,'// Row 0: T == 0'); f_fill=find(fast,'const uint32_t gapv  = pack2'); f_sink=find(fast,'// sink selection: first strict'); f_tb=find(fast,'// Second half: traceback')
R+= [('poa_fast.cuh',f_pred-2,f_fm-1,'pred_row'),('poa_fast.cuh',f_fm,f_lt-4,'tilted_score_global'),('poa_fast.cuh',f_lt-3,f_al-2,'load_tile'),('poa_fast.cuh',f_al-1,f_pre-1,'fast setup'),('poa_fast.cuh',f_pre,f_fill-1,'fast prepass+row0'),('poa_fast.cuh',f_fill,f_sink-1,'fast fill'),('poa_fast.cuh',f_sink,f_tb-1,'fast sink'),('poa_fast.cuh',f_tb,len(fast),'fast traceback')]
d_sm=find(dev,'explicit shared-memory accessors'); d_g=find(dev,'struct Graph'); d_bb=find(dev,'__device__ void init_backbone'); d_add=find(dev,'__device__ int32_t add_alignment'); d_kg=find(dev,'__device__ void topsort_kahn('); d_ks=find(dev,'__device__ bool topsort_kahn_smem'); d_sp=find(dev,'__device__ bool topsort_spoa_order'); d_bc=find(dev,'__device__ int32_t branch_completion'); d_gc=find(dev,'__device__ int32_t generate_consensus('); d_gs=find(dev,'__device__ int32_t generate_consensus_smem'); d_msa=find(dev,'__device__ int32_t msa_columns')
R+= [('poa_device.cuh',d_sm,d_g-2,'smem helpers'),('poa_device.cuh',d_g-1,d_g+60,'graph accessors'),('poa_device.cuh',d_bb-2,d_add-12,'backbone'),('poa_device.cuh',d_add-11,d_kg-3,'add_alignment'),('poa_device.cuh',d_kg-2,d_ks-10,'kahn global'),('poa_device.cuh',d_ks-9,d_sp-4,'kahn smem'),('poa_device.cuh',d_bc-3,d_gs-12,'consensus global'),('poa_device.cuh',d_gs-11,d_msa-3,'consensus smem')]
print(R)
