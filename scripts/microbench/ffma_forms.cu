#include <cuda_runtime.h>
#define H1 64
#define H2 64
#define D 2
struct alignas(16) NW { float w[D*H1 + H1 + H1*H2 + H2 + H2*D + D]; };
// layout offsets
#define OW1 0
#define OB1 (OW1 + H1*D)
#define OW2 (OB1 + H1)
#define OB2 (OW2 + H2*H1)
#define OW3 (OB2 + H2)
#define OB3 (OW3 + D*H2)
__device__ __forceinline__ float th(float x){ float x2=x*x; float n=fmaf(x2,fmaf(x2,fmaf(x2,fmaf(x2,1.587199e-8f,2.2332108e-5f),0.0035974074f),0.1346604f),1.0f);
 float d=fmaf(x2,fmaf(x2,fmaf(x2,fmaf(x2,8.7767893e-7f,0.0003453992f),0.026262015f),0.4679937f),1.0f); return x2<66.f? x*__fdividef(n,d): copysignf(1.f,x);}
template<int UNROLL_O>
__global__ void __launch_bounds__(256,1) kA(const __grid_constant__ NW p, const float* __restrict__ zin, float* __restrict__ out, int B, int Bs) {
  int b = blockIdx.x*blockDim.x+threadIdx.x; if (b>=B) return;
  float z[D]; for(int i=0;i<D;++i) z[i]=zin[i*Bs+b];
  float h1[H1], t1[D][H1];
  #pragma unroll
  for(int o=0;o<H1;++o){ float a=p.w[OB1+o];
    #pragma unroll
    for(int i=0;i<D;++i) a=fmaf(p.w[OW1+o*D+i], z[i], a);
    float h=th(a); h1[o]=h; float d=fmaf(-h,h,1.f);
    #pragma unroll
    for(int v=0;v<D;++v) t1[v][o]=p.w[OW1+o*D+v]*d; }
  float p3[D], r3[D][D];
  #pragma unroll
  for(int j=0;j<D;++j){ p3[j]=p.w[OB3+j];
    #pragma unroll
    for(int v=0;v<D;++v) r3[v][j]=0.f; }
  #pragma unroll UNROLL_O
  for(int o=0;o<H2;++o){
    float a=p.w[OB2+o]; float t[D];
    #pragma unroll
    for(int v=0;v<D;++v) t[v]=0.f;
    #pragma unroll
    for(int i=0;i<H1;++i){ float w=p.w[OW2+o*H1+i]; a=fmaf(w,h1[i],a);
      #pragma unroll
      for(int v=0;v<D;++v) t[v]=fmaf(w,t1[v][i],t[v]); }
    float h=th(a); float d=fmaf(-h,h,1.f);
    #pragma unroll
    for(int j=0;j<D;++j){ float w3=p.w[OW3+j*H2+o]; p3[j]=fmaf(w3,h,p3[j]);
      #pragma unroll
      for(int v=0;v<D;++v) r3[v][j]=fmaf(w3,t[v]*d,r3[v][j]); }
  }
  float tr=0.f;
  #pragma unroll
  for(int j=0;j<D;++j){ float y=th(p3[j]); out[j*Bs+b]=y; tr -= fmaf(-y,y,1.f)*r3[j][j]; }
  out[D*Bs+b]=tr;
}

// ---- variant S: weights of layer 2 in shared memory (broadcast LDS.128), rest from params
__global__ void __launch_bounds__(256,1) kS(const __grid_constant__ NW p, const float* __restrict__ zin, float* __restrict__ out, int B, int Bs) {
  __shared__ __align__(16) float sW2[H2*H1];
  for (int i=threadIdx.x;i<H2*H1;i+=blockDim.x) sW2[i]=p.w[OW2+i];
  __syncthreads();
  int b = blockIdx.x*blockDim.x+threadIdx.x; if (b>=B) return;
  float z[D]; for(int i=0;i<D;++i) z[i]=zin[i*Bs+b];
  float h1[H1], t1[D][H1];
  #pragma unroll
  for(int o=0;o<H1;++o){ float a=p.w[OB1+o];
    #pragma unroll
    for(int i=0;i<D;++i) a=fmaf(p.w[OW1+o*D+i], z[i], a);
    float h=th(a); h1[o]=h; float d=fmaf(-h,h,1.f);
    #pragma unroll
    for(int v=0;v<D;++v) t1[v][o]=p.w[OW1+o*D+v]*d; }
  float p3[D], r3[D][D];
  #pragma unroll
  for(int j=0;j<D;++j){ p3[j]=p.w[OB3+j];
    #pragma unroll
    for(int v=0;v<D;++v) r3[v][j]=0.f; }
  #pragma unroll 2
  for(int o=0;o<H2;++o){
    float a=p.w[OB2+o]; float t[D];
    #pragma unroll
    for(int v=0;v<D;++v) t[v]=0.f;
    const float4* wr = reinterpret_cast<const float4*>(sW2 + o*H1);
    #pragma unroll
    for(int i4=0;i4<H1/4;++i4){ float4 w4=wr[i4]; float ww[4]={w4.x,w4.y,w4.z,w4.w};
      #pragma unroll
      for(int k=0;k<4;++k){ int i=i4*4+k; a=fmaf(ww[k],h1[i],a);
        #pragma unroll
        for(int v=0;v<D;++v) t[v]=fmaf(ww[k],t1[v][i],t[v]); } }
    float h=th(a); float d=fmaf(-h,h,1.f);
    #pragma unroll
    for(int j=0;j<D;++j){ float w3=p.w[OW3+j*H2+o]; p3[j]=fmaf(w3,h,p3[j]);
      #pragma unroll
      for(int v=0;v<D;++v) r3[v][j]=fmaf(w3,t[v]*d,r3[v][j]); }
  }
  float tr=0.f;
  #pragma unroll
  for(int j=0;j<D;++j){ float y=th(p3[j]); out[j*Bs+b]=y; tr -= fmaf(-y,y,1.f)*r3[j][j]; }
  out[D*Bs+b]=tr;
}
// ---- FFMA operand-form probes: dependent chains, 8 independent accumulators
template<int FORM> __global__ void kF(const __grid_constant__ NW p, float* out, int iters) {
  float a[8]; for(int k=0;k<8;++k) a[k]=threadIdx.x*1e-3f+k;
  float r0=out[0], r1=out[1];
  for(int it=0;it<iters;++it){
    #pragma unroll
    for(int u=0;u<64;++u){
      #pragma unroll
      for(int k=0;k<8;++k){
        if(FORM==0) a[k]=fmaf(a[k],0.999f,1e-3f);          // imm
        else if(FORM==1) a[k]=fmaf(a[k],p.w[u],a[k]);      // const/UR operand + 2 regs (acc=acc*w+acc)
        else if(FORM==2) a[k]=fmaf(a[k],r0,r1);            // 3 distinct regs
        else a[k]=fmaf(p.w[u],r0,a[k]);                    // w(UR)*reg + acc : the MLP form
      }
    }
  }
  float s=0; for(int k=0;k<8;++k) s+=a[k]; out[blockIdx.x*blockDim.x+threadIdx.x]=s;
}
#include <cstdio>
#include <vector>
template<class F> float timeit(F f, int reps){ cudaEvent_t a,b; cudaEventCreate(&a); cudaEventCreate(&b); f(); cudaDeviceSynchronize(); cudaEventRecord(a); for(int i=0;i<reps;++i) f(); cudaEventRecord(b); cudaEventSynchronize(b); float ms; cudaEventElapsedTime(&ms,a,b); return ms/reps; }
int main(){
  const int B=1<<20, Bs=B; NW hw; for(size_t i=0;i<sizeof(hw.w)/4;++i) hw.w[i]=0.1f*sinf(0.37f*i);
  float *zin,*out; cudaMalloc(&zin,sizeof(float)*D*Bs); cudaMalloc(&out,sizeof(float)*(D+1)*Bs);
  std::vector<float> hz(D*Bs); for(size_t i=0;i<hz.size();++i) hz[i]=sinf(0.001f*i); cudaMemcpy(zin,hz.data(),hz.size()*4,cudaMemcpyHostToDevice);
  const double fl = 26112.0*B;
  float ms;
  ms=timeit([&]{ kA<1><<<B/256,256>>>(hw,zin,out,B,Bs); },10); printf("kA<1>  %.3f ms  %.2f TFLOP/s\n",ms,fl/ms/1e9);
  ms=timeit([&]{ kA<2><<<B/256,256>>>(hw,zin,out,B,Bs); },10); printf("kA<2>  %.3f ms  %.2f TFLOP/s\n",ms,fl/ms/1e9);
  ms=timeit([&]{ kA<4><<<B/256,256>>>(hw,zin,out,B,Bs); },10); printf("kA<4>  %.3f ms  %.2f TFLOP/s\n",ms,fl/ms/1e9);
  ms=timeit([&]{ kA<8><<<B/256,256>>>(hw,zin,out,B,Bs); },10); printf("kA<8>  %.3f ms  %.2f TFLOP/s\n",ms,fl/ms/1e9);
  ms=timeit([&]{ kA<64><<<B/256,256>>>(hw,zin,out,B,Bs); },10); printf("kA<64> %.3f ms  %.2f TFLOP/s\n",ms,fl/ms/1e9);
  ms=timeit([&]{ kS<<<B/256,256>>>(hw,zin,out,B,Bs); },10); printf("kS     %.3f ms  %.2f TFLOP/s\n",ms,fl/ms/1e9);
  ms=timeit([&]{ kA<4><<<B/128,128>>>(hw,zin,out,B,Bs); },10); printf("kA<4> 128thr %.3f ms  %.2f TFLOP/s\n",ms,fl/ms/1e9);
  const int blocks=148*8, iters=512; const double ff=2.0*8*64*iters*(double)blocks*256;
  ms=timeit([&]{ kF<0><<<blocks,256>>>(hw,out,iters); },5); printf("FFMA imm      %.2f TFLOP/s\n",ff/ms/1e9);
  ms=timeit([&]{ kF<1><<<blocks,256>>>(hw,out,iters); },5); printf("FFMA R,UR,R   %.2f TFLOP/s\n",ff/ms/1e9);
  ms=timeit([&]{ kF<2><<<blocks,256>>>(hw,out,iters); },5); printf("FFMA R,R,R    %.2f TFLOP/s\n",ff/ms/1e9);
  ms=timeit([&]{ kF<3><<<blocks,256>>>(hw,out,iters); },5); printf("FFMA UR,R,acc %.2f TFLOP/s\n",ff/ms/1e9);
  printf("%s\n", cudaGetErrorString(cudaGetLastError()));
  return 0;
}
