// Random-access roof of this GPU's HBM: how many scattered 32-byte sectors per second can the memory system deliver?
// Each thread issues UNROLL independent 16-byte loads (one sector each) at hashed places of a table far larger than L2, many
// times; one xor result per thread is stored so that nothing is optimised away. Build: nvcc -O3 -gencode
// arch=compute_100a,code=sm_100a -o tools/_build/random_sectors tools/random_sectors.cu ; prints one JSON line.
// Why: at chromosome scale the rewrite kernel touches ~16 sectors per record in different DRAM pages;
// MEASURED_PEAKS.json:hbm_gbs is the roof for streamed bytes, this is the roof for scattered sectors.
#include <cstdint>
#include <cstdio>
#include <cuda_runtime.h>

__device__ __forceinline__ uint64_t mix(uint64_t x)
{
	x ^= x >> 33; x *= 0xff51afd7ed558ccdull; x ^= x >> 33; x *= 0xc4ceb9fe1a85ec53ull; x ^= x >> 33;
	return x;
}

template <int UNROLL, int SECTORS>     // SECTORS consecutive sectors per gather (1 = 32 B, 2 = 64 B, 4 = 128 B)
__global__ void gather_kernel(const uint4 *table, uint64_t n_rows, uint32_t rounds, uint64_t seed, uint32_t *out)
{
	uint64_t const tid = (uint64_t) blockIdx.x * blockDim.x + threadIdx.x;
	uint32_t acc = 0;
	for (uint32_t r = 0; r < rounds; ++r) {
		uint4 v[UNROLL][SECTORS];
		#pragma unroll
		for (int u = 0; u < UNROLL; ++u) {
			uint64_t const row = mix(seed + (tid * rounds + r) * UNROLL + u) % n_rows;
			#pragma unroll
			for (int s = 0; s < SECTORS; ++s) v[u][s] = __ldg(table + (row * SECTORS + s) * 2);      // first 16 bytes of each 32-byte sector
		}
		#pragma unroll
		for (int u = 0; u < UNROLL; ++u)
			#pragma unroll
			for (int s = 0; s < SECTORS; ++s) acc ^= v[u][s].x ^ v[u][s].w;
	}
	out[tid] = acc;
}

template <int UNROLL, int SECTORS>
static double run(const uint4 *table, uint64_t table_bytes, uint32_t *out, int blocks, int threads, uint32_t rounds)
{
	uint64_t const n_rows = table_bytes / (32ull * SECTORS);
	cudaEvent_t a, b; cudaEventCreate(&a); cudaEventCreate(&b);
	gather_kernel <UNROLL, SECTORS><<<blocks, threads>>>(table, n_rows, rounds, 1, out);
	float best = 1e30f;
	for (int i = 0; i < 5; ++i) {
		cudaEventRecord(a);
		gather_kernel <UNROLL, SECTORS><<<blocks, threads>>>(table, n_rows, rounds, 17 + i, out);
		cudaEventRecord(b); cudaEventSynchronize(b);
		float ms = 0; cudaEventElapsedTime(&ms, a, b);
		if (ms < best) best = ms;
	}
	double const gathers = (double) blocks * threads * rounds * UNROLL;
	return gathers / (best * 1e-3);
}

int main()
{
	uint64_t const table_bytes = 8ull << 30;
	uint4 *table = nullptr; uint32_t *out = nullptr;
	if (cudaMalloc(&table, table_bytes) != cudaSuccess) { fprintf(stderr, "cudaMalloc failed\n"); return 1; }
	cudaMemset(table, 1, table_bytes);
	int sms = 148; cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, 0);
	int const threads = 256, blocks = sms * 8;
	cudaMalloc(&out, (size_t) blocks * threads * 4);
	uint32_t const rounds = 64;
	double const s1_u4 = run <4, 1>(table, table_bytes, out, blocks, threads, rounds);
	double const s1_u8 = run <8, 1>(table, table_bytes, out, blocks, threads, rounds);
	double const s1_u16 = run <16, 1>(table, table_bytes, out, blocks, threads, rounds / 2);
	double const s2 = run <8, 2>(table, table_bytes, out, blocks, threads, rounds / 2);
	double const s4 = run <4, 4>(table, table_bytes, out, blocks, threads, rounds / 2);
	cudaError_t const e = cudaDeviceSynchronize();
	printf("{\"table_bytes\": %llu, \"sms\": %d, \"error\": \"%s\", \"gathers_32B_per_s\": {\"4_in_flight_per_thread\": %.4g, \"8_in_flight\": %.4g, \"16_in_flight\": %.4g}, "
	       "\"gathers_64B_per_s\": %.4g, \"gathers_128B_per_s\": %.4g, \"sectors_per_s_at_128B\": %.4g}\n",
	       (unsigned long long) table_bytes, sms, cudaGetErrorString(e), s1_u4, s1_u8, s1_u16, s2, s4, 4 * s4);
	return 0;
}
