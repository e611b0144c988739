/*
 * pipe_rates.cu -- issue-rate probe for the instruction kinds the shadow walk is
 * built from (sm_100a).  Each kernel runs 8 independent dependency chains of one
 * instruction kind per thread; the result is warp instructions per clock per SM,
 * which tells which pipe a kind shares and at what rate (B300_MICROARCH.md: fma
 * and alu pipes, one warp instruction per two clocks per scheduler each).
 *
 *   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o build/pipe_rates tools/pipe_rates.cu
 */
#include <cuda_fp16.h>
#include <cstdint>
#include <cstdio>
#include <cuda_runtime.h>

#define CHAINS 8
#define UNROLL 8

template <int KIND>
__device__ __forceinline__ uint32_t step(uint32_t v, uint32_t a, uint32_t b)
{
	uint32_t r;
	if (KIND == 0) { /* FFMA */
		asm volatile("fma.rn.f32 %0, %1, %2, %3;" : "=r"(r) : "r"(v), "r"(a), "r"(b));
	} else if (KIND == 1) { /* FMNMX (two per step, min then max: ptxas fuses max(max()) into FMNMX3) */
		uint32_t t;
		asm volatile("min.f32 %0, %1, %2;" : "=r"(t) : "r"(v), "r"(b));
		asm volatile("max.f32 %0, %1, %2;" : "=r"(r) : "r"(t), "r"(a));
	} else if (KIND == 2) { /* FMNMX3 */
		asm volatile("max.f32 %0, %1, %2, %3;" : "=r"(r) : "r"(v), "r"(a), "r"(b));
	} else if (KIND == 3) { /* PRMT */
		asm volatile("prmt.b32 %0, %1, %2, %3;" : "=r"(r) : "r"(v), "r"(a), "r"(b));
	} else if (KIND == 4) { /* HFMA2 */
		asm volatile("fma.rn.f16x2 %0, %1, %2, %3;" : "=r"(r) : "r"(v), "r"(a), "r"(b));
	} else if (KIND == 5) { /* HADD2 (sub) */
		asm volatile("sub.rn.f16x2 %0, %1, %2;" : "=r"(r) : "r"(v), "r"(a));
	} else if (KIND == 6) { /* HMUL2 */
		asm volatile("mul.rn.f16x2 %0, %1, %2;" : "=r"(r) : "r"(v), "r"(a));
	} else if (KIND == 7) { /* HMNMX2 (two per step, min then max) */
		uint32_t t;
		asm volatile("min.f16x2 %0, %1, %2;" : "=r"(t) : "r"(v), "r"(b));
		asm volatile("max.f16x2 %0, %1, %2;" : "=r"(r) : "r"(t), "r"(a));
	} else if (KIND == 8) { /* 3-input half2 max as the compiler fuses it (VHMNMX) */
		__half2 x = *reinterpret_cast<__half2 *>(&v), y = *reinterpret_cast<__half2 *>(&a), z = *reinterpret_cast<__half2 *>(&b);
		__half2 w = __hmax2(__hmax2(x, y), z);
		r = *reinterpret_cast<uint32_t *>(&w);
	} else if (KIND == 9) { /* HSET2 mask */
		asm volatile("set.le.u32.f16x2 %0, %1, %2;" : "=r"(r) : "r"(v), "r"(a));
	} else if (KIND == 10) { /* F2FP pack */
		asm volatile("cvt.rn.f16x2.f32 %0, %1, %2;" : "=r"(r) : "r"(v), "r"(a));
	} else if (KIND == 11) { /* LOP3 */
		asm volatile("lop3.b32 %0, %1, %2, %3, 0x96;" : "=r"(r) : "r"(v), "r"(a), "r"(b));
	} else if (KIND == 12) { /* IMAD */
		asm volatile("mad.lo.u32 %0, %1, %2, %3;" : "=r"(r) : "r"(v), "r"(a), "r"(b));
	} else if (KIND == 13) { /* FSEL via setp+selp */
		asm volatile("{ .reg .pred p; setp.lt.f32 p, %1, %2; selp.b32 %0, %1, %3, p; }" : "=r"(r) : "r"(v), "r"(a), "r"(b));
	} else if (KIND == 14) { /* FADD */
		asm volatile("add.rn.f32 %0, %1, %2;" : "=r"(r) : "r"(v), "r"(a));
	} else if (KIND == 15) { /* mixed: one FFMA + one FMNMX per step (do the two pipes overlap?) */
		uint32_t t;
		asm volatile("fma.rn.f32 %0, %1, %2, %3;" : "=r"(t) : "r"(v), "r"(a), "r"(b));
		asm volatile("max.f32 %0, %1, %2;" : "=r"(r) : "r"(t), "r"(a));
	} else if (KIND == 16) { /* mixed: HMUL2 + HMNMX2 */
		uint32_t t;
		asm volatile("mul.rn.f16x2 %0, %1, %2;" : "=r"(t) : "r"(v), "r"(a));
		asm volatile("max.f16x2 %0, %1, %2;" : "=r"(r) : "r"(t), "r"(b));
	} else { /* IADD3 + LOP3 pair (an add chain alone is folded into IMAD/LEA) */
		uint32_t t;
		asm volatile("add.u32 %0, %1, %2;" : "=r"(t) : "r"(v), "r"(a));
		asm volatile("lop3.b32 %0, %1, %2, %3, 0x96;" : "=r"(r) : "r"(t), "r"(a), "r"(b));
	}
	return r;
}

template <int KIND>
__global__ void __launch_bounds__(256) rate_kernel(int iters, uint32_t a, uint32_t b, uint32_t *sink)
{
	uint32_t v[CHAINS];
#pragma unroll
	for (int c = 0; c < CHAINS; c++)
		v[c] = threadIdx.x * 2654435761u + c;
	for (int i = 0; i < iters; i++) {
#pragma unroll
		for (int u = 0; u < UNROLL; u++)
#pragma unroll
			for (int c = 0; c < CHAINS; c++)
				v[c] = step<KIND>(v[c], a, b);
	}
	uint32_t s = 0;
#pragma unroll
	for (int c = 0; c < CHAINS; c++)
		s ^= v[c];
	if (s == 0x12345678u)
		sink[0] = s;
}

template <int KIND>
void run(const char *name, int per_step, int sms, double clock_ghz, uint32_t *sink)
{
	const int iters = 2048, blocks = sms * 8, threads = 256;
	cudaEvent_t e0, e1;
	cudaEventCreate(&e0);
	cudaEventCreate(&e1);
	float best = 1e30f;
	for (int rep = 0; rep < 4; rep++) {
		cudaEventRecord(e0);
		rate_kernel<KIND><<<blocks, threads>>>(iters, 0x3c003c00u, 0x3f800000u, sink);
		cudaEventRecord(e1);
		cudaEventSynchronize(e1);
		float ms;
		cudaEventElapsedTime(&ms, e0, e1);
		if (rep && ms < best)
			best = ms;
	}
	const double warp_instr = (double)iters * UNROLL * CHAINS * per_step * blocks * (threads / 32);
	const double per_clk_sm = warp_instr / (best * 1e-3) / (clock_ghz * 1e9) / sms;
	printf("%-28s %8.3f ms  %6.3f warp-instr/clk/SM  (%5.3f per scheduler)\n", name, best, per_clk_sm, per_clk_sm / 4);
	cudaEventDestroy(e0);
	cudaEventDestroy(e1);
}

int main()
{
	int sms = 0, khz = 0;
	cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, 0);
	cudaDeviceGetAttribute(&khz, cudaDevAttrClockRate, 0);
	const double ghz = khz * 1e-6;
	uint32_t *sink;
	cudaMalloc(&sink, 4);
	printf("SMs %d, clock %.3f GHz (rates assume the kernel runs at this clock)\n", sms, ghz);
	run<0>("FFMA", 1, sms, ghz, sink);
	run<14>("FADD", 1, sms, ghz, sink);
	run<1>("FMNMX (min,max)", 2, sms, ghz, sink);
	run<2>("FMNMX3", 1, sms, ghz, sink);
	run<3>("PRMT", 1, sms, ghz, sink);
	run<11>("LOP3", 1, sms, ghz, sink);
	run<17>("IADD3+LOP3", 2, sms, ghz, sink);
	run<12>("IMAD", 1, sms, ghz, sink);
	run<13>("FSETP+SEL", 2, sms, ghz, sink);
	run<4>("HFMA2", 1, sms, ghz, sink);
	run<5>("HADD2 (sub.f16x2)", 1, sms, ghz, sink);
	run<6>("HMUL2", 1, sms, ghz, sink);
	run<7>("HMNMX2 (min,max)", 2, sms, ghz, sink);
	run<8>("half2 max3 (VHMNMX?)", 1, sms, ghz, sink);
	run<9>("HSET2 (set.le.u32.f16x2)", 1, sms, ghz, sink);
	run<10>("F2FP.PACK_AB", 1, sms, ghz, sink);
	run<15>("FFMA+FMNMX interleaved", 2, sms, ghz, sink);
	run<16>("HMUL2+HMNMX2 interleaved", 2, sms, ghz, sink);
	if (cudaDeviceSynchronize() != cudaSuccess) {
		printf("CUDA error: %s\n", cudaGetErrorString(cudaGetLastError()));
		return 1;
	}
	return 0;
}
