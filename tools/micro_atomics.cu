// Micro-benchmark (design evidence, not product): L2-resident keyed-update cost on B200.
// Build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -lineinfo -o tools/micro_atomics tools/micro_atomics.cu
// Each variant streams n events (ts,key,val = 24 B) and updates a slot table of `slots` x 32 B.
#include <cstdio>
#include <cstdint>
#include <cstdlib>
#include <cuda_runtime.h>
#define CK(x) do{cudaError_t e=(x); if(e!=cudaSuccess){printf("CUDA %s @%d\n",cudaGetErrorString(e),__LINE__);exit(1);} }while(0)

__device__ __forceinline__ uint64_t mix64(uint64_t x){ x^=x>>30; x*=0xbf58476d1ce4e5b9ULL; x^=x>>27; x*=0x94d049bb133111ebULL; x^=x>>31; return x; }

__global__ void gen(int64_t* ts, uint64_t* key, double* val, size_t n, uint64_t nkeys){
  size_t i = blockIdx.x*(size_t)blockDim.x+threadIdx.x; size_t st=(size_t)gridDim.x*blockDim.x;
  for(;i<n;i+=st){ uint64_t h=mix64(i*0x9E3779B97F4A7C15ULL+42); ts[i]=(int64_t)(i/1000); key[i]=h%nkeys; val[i]=(double)(mix64(h)>>11)*(100.0/9007199254740992.0);} }

struct Slot { unsigned long long count; double sum; unsigned long long mx; unsigned long long mn; };

__device__ __forceinline__ unsigned long long enc(double v){ unsigned long long b=__double_as_longlong(v); return (b>>63)? ~b : (b|0x8000000000000000ULL); }

// variant 0: read only (sum into a sink to keep loads)
// variant 1: +1 red.add.u64 (count)
// variant 2: count + sum(f64) + max(u64)  : 3 separate red instructions
// variant 3: 4 fields, 4 separate reds
// variant 4: pair-lane: two lanes cooperate per event; lane parity picks the field: instr A = red.add.f64 {countf,sum}; instr B = red.max.u64 {mx, ~mn}
// variant 5: as 2 but with a directory probe (dependent 8B load of keys[slot]) first
// variant 6: as 4 but with the probe
template<int V>
__global__ void __launch_bounds__(256) upd(const int64_t* __restrict__ ts, const uint64_t* __restrict__ key, const double* __restrict__ val,
     size_t n, Slot* tab, const uint64_t* __restrict__ dir, uint64_t mask, unsigned long long* sink){
  size_t i = blockIdx.x*(size_t)blockDim.x+threadIdx.x; size_t st=(size_t)gridDim.x*blockDim.x;
  unsigned long long acc=0;
  for(;i<n;i+=st){
    int64_t t=ts[i]; uint64_t k=key[i]; double v=val[i];
    uint64_t s = mix64(k)&mask;
    if(V==5||V==6){ uint64_t dk=dir[s]; if(dk!=k) s=(s+1)&mask; }
    if(V==0){ acc+= (unsigned long long)t + k + __double_as_longlong(v); }
    if(V==1){ atomicAdd(&tab[s].count,1ULL); }
    if(V==2||V==5){ atomicAdd(&tab[s].count,1ULL); atomicAdd(&tab[s].sum,v); atomicMax(&tab[s].mx,enc(v)); }
    if(V==3){ atomicAdd(&tab[s].count,1ULL); atomicAdd(&tab[s].sum,v); atomicMax(&tab[s].mx,enc(v)); atomicMax(&tab[s].mn,~enc(v)); }
    if(V==4||V==6){
      // exchange with partner lane so that each instruction carries both fields of an event in adjacent lanes
      unsigned lane=threadIdx.x&31; bool odd=lane&1;
      uint64_t s_p=__shfl_xor_sync(0xffffffffu,s,1); double v_p=__shfl_xor_sync(0xffffffffu,v,1);
      // round 0 handles even lane's event, round 1 the odd lane's event
      #pragma unroll
      for(int r=0;r<2;r++){
        uint64_t ss = (r==(int)odd)? s : s_p; double vv=(r==(int)odd)? v : v_p;
        double* pa = odd? &tab[ss].sum : reinterpret_cast<double*>(&tab[ss].count);
        atomicAdd(pa, odd? vv : 1.0);
        unsigned long long* pm = odd? &tab[ss].mn : &tab[ss].mx;
        atomicMax(pm, odd? ~enc(vv) : enc(vv));
      }
    }
  }
  if(V==0 && acc==0x1234567) *sink=acc;
}

template<int V> float run(const int64_t* ts,const uint64_t* key,const double* val,size_t n,Slot* tab,uint64_t* dir,uint64_t slots,unsigned long long* sink,int blocks){
  cudaEvent_t a,b; CK(cudaEventCreate(&a)); CK(cudaEventCreate(&b));
  float best=1e30f;
  for(int it=0;it<5;it++){
    CK(cudaMemsetAsync(tab,0,slots*sizeof(Slot)));
    CK(cudaEventRecord(a));
    upd<V><<<blocks,256>>>(ts,key,val,n,tab,dir,slots-1,sink);
    CK(cudaEventRecord(b)); CK(cudaEventSynchronize(b)); float ms; CK(cudaEventElapsedTime(&ms,a,b)); if(it>0&&ms<best)best=ms;
  }
  CK(cudaGetLastError());
  return best;
}
__global__ void filldir(uint64_t* dir, uint64_t slots, uint64_t nkeys){ size_t i=blockIdx.x*(size_t)blockDim.x+threadIdx.x; if(i<nkeys){ dir[mix64(i)&(slots-1)]=i; } }

int main(int argc,char**argv){
  size_t n = (argc>1)? strtoull(argv[1],0,10) : (1ull<<26);
  uint64_t nkeys = (argc>2)? strtoull(argv[2],0,10) : 1000000ull;
  uint64_t slots=1; while(slots<2*nkeys) slots<<=1;
  int64_t* ts; uint64_t* key; double* val; Slot* tab; uint64_t* dir; unsigned long long* sink;
  CK(cudaMalloc(&ts,n*8)); CK(cudaMalloc(&key,n*8)); CK(cudaMalloc(&val,n*8)); CK(cudaMalloc(&tab,slots*sizeof(Slot))); CK(cudaMalloc(&dir,slots*8)); CK(cudaMalloc(&sink,8));
  CK(cudaMemset(dir,0xff,slots*8));
  gen<<<148*8,256>>>(ts,key,val,n,nkeys); filldir<<<(nkeys+255)/256,256>>>(dir,slots,nkeys); CK(cudaDeviceSynchronize());
  printf("n=%zu keys=%llu slots=%llu table=%.1f MB\n",n,(unsigned long long)nkeys,(unsigned long long)slots,slots*32/1e6);
  for(int bm=8; bm<=16; bm*=2){
    int blocks=148*bm;
    float t0=run<0>(ts,key,val,n,tab,dir,slots,sink,blocks);
    float t1=run<1>(ts,key,val,n,tab,dir,slots,sink,blocks);
    float t2=run<2>(ts,key,val,n,tab,dir,slots,sink,blocks);
    float t3=run<3>(ts,key,val,n,tab,dir,slots,sink,blocks);
    float t4=run<4>(ts,key,val,n,tab,dir,slots,sink,blocks);
    float t5=run<5>(ts,key,val,n,tab,dir,slots,sink,blocks);
    float t6=run<6>(ts,key,val,n,tab,dir,slots,sink,blocks);
    const char* nm[7]={"read-only","1 red","3 red","4 red","pair-lane 4 fields","probe+3 red","probe+pair-lane"};
    float t[7]={t0,t1,t2,t3,t4,t5,t6};
    for(int v=0;v<7;v++) printf("blocks=%d %-20s %8.3f ms  %7.2f Gev/s  %7.1f GB/s(24B/ev)\n",blocks,nm[v],t[v],n/t[v]/1e6,24.0*n/t[v]/1e6);
  }
  return 0;
}
