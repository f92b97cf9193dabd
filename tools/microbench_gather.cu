// Developer microbenchmark: how fast can ONE SM-resident CTA per SM gather 3 KB rows with
// cp.async.bulk, as a function of the address pattern?  (Explains the sparse-mask behaviour.)
#include <cstdio>
#include <cstdint>
#include <cstdlib>
#include <vector>
#include <cuda_runtime.h>
#define CK(x) do{cudaError_t e=(x); if(e!=cudaSuccess){printf("%s: %s\n",#x,cudaGetErrorString(e)); exit(1);} }while(0)
__device__ __forceinline__ uint32_t s32(const void*p){return (uint32_t)__cvta_generic_to_shared(p);}
__global__ void __launch_bounds__(128,1) gather(const unsigned char* E, const uint32_t* rows, uint32_t per_cta, uint32_t row_bytes, uint32_t inflight, unsigned long long* cycles) {
    extern __shared__ __align__(128) unsigned char smem[];
    __shared__ uint64_t bars[64];
    if (threadIdx.x==0){ for(int i=0;i<64;i++) asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;"::"r"(s32(&bars[i]))); asm volatile("fence.mbarrier_init.release.cluster;":::"memory"); }
    __syncthreads();
    if (threadIdx.x==0) {
        long long t0=clock64();
        const uint32_t* my = rows + (size_t)blockIdx.x*per_cta;
        uint32_t issued=0, done=0;
        while (done<per_cta) {
            while (issued<per_cta && issued-done<inflight) {
                uint32_t s=issued%inflight; uint32_t bar=s32(&bars[s]);
                asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;"::"r"(bar),"r"(row_bytes):"memory");
                asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"::"r"(s32(smem+(size_t)s*row_bytes)),"l"(E+(size_t)my[issued]*row_bytes),"r"(row_bytes),"r"(bar):"memory");
                issued++;
            }
            uint32_t s=done%inflight; uint32_t bar=s32(&bars[s]); uint32_t par=(done/inflight)&1, ok;
            do { asm volatile("{.reg .pred p; mbarrier.try_wait.parity.shared::cta.b64 p,[%1],%2; selp.u32 %0,1,0,p;}":"=r"(ok):"r"(bar),"r"(par):"memory"); } while(!ok);
            done++;
        }
        cycles[blockIdx.x]=clock64()-t0;
    }
}
int main(int argc,char**argv){
    const uint32_t row_bytes=3072; const uint64_t n=1<<20; int sms=148;
    unsigned char* E; CK(cudaMalloc(&E,n*row_bytes)); CK(cudaMemset(E,1,n*row_bytes));
    CK(cudaFuncSetAttribute(gather,cudaFuncAttributeMaxDynamicSharedMemorySize,200*1024));
    unsigned long long* cyc; CK(cudaMalloc(&cyc,sms*8));
    const uint32_t per=512;
    for (int pattern=0; pattern<4; pattern++) for (uint32_t inflight : {1u,4u,16u,64u}) {
        std::vector<uint32_t> rows((size_t)sms*per);
        uint64_t st=88172645463325252ull;
        for (int c=0;c<sms;c++) for(uint32_t i=0;i<per;i++){ st^=st<<13; st^=st>>7; st^=st<<17;
            uint32_t r;
            if(pattern==0) r=(uint32_t)(st%n);                                   // random over 3.2 GB
            else if(pattern==1) r=(uint32_t)(((uint64_t)c*7000+ (st%680))%n);     // random within a ~2 MB window per CTA
            else if(pattern==2) r=(uint32_t)(((uint64_t)c*7000+i)%n);             // consecutive rows
            else r=(uint32_t)(((uint64_t)c*7000+ (uint64_t)i*8)%n);              // every 8th row (12.5% mask, regular)
            rows[(size_t)c*per+i]=r; }
        uint32_t* d; CK(cudaMalloc(&d,rows.size()*4)); CK(cudaMemcpy(d,rows.data(),rows.size()*4,cudaMemcpyHostToDevice));
        cudaEvent_t e0,e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
        gather<<<sms,128,inflight*row_bytes>>>(E,d,per,row_bytes,inflight,cyc); CK(cudaDeviceSynchronize());
        cudaEventRecord(e0); gather<<<sms,128,inflight*row_bytes>>>(E,d,per,row_bytes,inflight,cyc); cudaEventRecord(e1); CK(cudaDeviceSynchronize());
        float ms; cudaEventElapsedTime(&ms,e0,e1);
        double gb=(double)sms*per*row_bytes/1e9;
        printf("pattern %d inflight %2u: %8.1f us  %7.1f GB/s  %6.2f us/row/SM\n",pattern,inflight,ms*1e3,gb/(ms/1e3),ms*1e3/per);
        cudaFree(d);
    }
    return 0;
}
