// f32x2_sel_probe.cu -- does ptxas fold half swaps / broadcasts / per-half negation into the operand selectors of
// FMUL2 / FFMA2 (as it does for FADD2)?  Build: nvcc -gencode arch=compute_100a,code=sm_100a -cubin, read with cuobjdump -sass.
typedef unsigned long long u64;
__device__ __forceinline__ u64 pk(float a, float b) { u64 r; asm("mov.b64 %0, {%1, %2};" : "=l"(r) : "f"(a), "f"(b)); return r; }
__device__ __forceinline__ void up(u64 v, float &a, float &b) { asm("mov.b64 {%0, %1}, %2;" : "=f"(a), "=f"(b) : "l"(v)); }
__device__ __forceinline__ u64 mul2(u64 a, u64 b) { u64 d; asm("mul.rn.f32x2 %0, %1, %2;" : "=l"(d) : "l"(a), "l"(b)); return d; }
__device__ __forceinline__ u64 fma2(u64 a, u64 b, u64 c) { u64 d; asm("fma.rn.f32x2 %0, %1, %2, %3;" : "=l"(d) : "l"(a), "l"(b), "l"(c)); return d; }

// complex multiply a * w with a = (re, im), w = (wr, wi), both as register pairs
__global__ void cmul_pairs(const u64 *a, const u64 *w, u64 *o, int n)
{
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    float ar, ai, wr, wi;
    up(a[i], ar, ai);
    up(w[i], wr, wi);
    const u64 t = mul2(a[i], pk(wr, wr));
    o[i] = fma2(pk(-ai, ar), pk(wi, wi), t);
}
// the same with a compile-time twiddle
__global__ void cmul_const(const u64 *a, u64 *o, int n)
{
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    float ar, ai;
    up(a[i], ar, ai);
    const u64 t = mul2(a[i], pk(0.92387953251128674f, 0.92387953251128674f));
    o[i] = fma2(pk(-ai, ar), pk(0.38268343236508977f, 0.38268343236508977f), t);
}
