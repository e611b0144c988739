/*
 * crt_rng.h -- the counter-based random stream shared by the CUDA path, the
 * CPU oracle and the instrumented reference build.
 *
 * The reference draws every random number from glibc rand() in program order
 * (rand_flt(), reference src/core/system.c:93-96), which no parallel renderer
 * can reproduce.  This header replaces that one function with a pure function
 * of a STRUCTURAL key, so that every implementation produces the same draw
 * for the same place in the reference's recursion:
 *
 *   u = crt_rand(node, c)
 *
 * node : 32-bit id of one cast_ray() invocation (reference render.c:136).
 *        The primary ray of pixel p has node = crt_node_root(seed, p); the
 *        k-th cast_ray() call made BY an invocation (its GI children first,
 *        render.c:285, then the reflection child :315, then the refraction
 *        child :338 -- whichever of them are actually made) has
 *        node = crt_node_child(parent, k).
 * c    : index of the draw inside that invocation, in program order:
 *        two draws per light sample (object.c:298-299 / :408) for every
 *        emitter != the hit object, in emitter order (render.c:170-178),
 *        then two draws per GI sample (render.c:281-282).
 *
 * Both values are known locally on the GPU (c from the node's own emitter
 * list, the child ordinal from the node's own spawn decisions), and on the
 * reference side they fall out of a per-invocation counter maintained by
 * gcc's -finstrument-functions hooks (oracle/ref_shim.c), so the reference
 * sources are compiled unmodified.
 *
 * Output is in [0,1) with 24 bits; the reference's rand_flt() is in [0,1]
 * inclusive -- irrelevant once both sides share this function.
 */
#ifndef CRT_RNG_H
#define CRT_RNG_H

#include <stdint.h>

#if defined(__CUDACC__)
#define CRT_RNG_FN __host__ __device__ static __forceinline__
#else
#define CRT_RNG_FN static inline
#endif

/* 32-bit finaliser ("lowbias32", public-domain constants by C. Wellons). */
CRT_RNG_FN uint32_t crt_mix32(uint32_t x)
{
	x ^= x >> 16;
	x *= 0x7feb352dU;
	x ^= x >> 15;
	x *= 0x846ca68bU;
	x ^= x >> 16;
	return x;
}

CRT_RNG_FN uint32_t crt_node_root(uint32_t seed, uint32_t pixel)
{
	return crt_mix32(pixel + 0x9e3779b9U * crt_mix32(seed ^ 0x5bd1e995U));
}

CRT_RNG_FN uint32_t crt_node_child(uint32_t parent, uint32_t ordinal)
{
	return crt_mix32(parent ^ (0x85ebca6bU * (ordinal + 1U)));
}

CRT_RNG_FN float crt_rand(uint32_t node, uint32_t c)
{
	uint32_t h = crt_mix32(node + 0xc2b2ae35U * (c + 1U));
	h = crt_mix32(h ^ node);
	return (float)(h >> 8) * (1.0f / 16777216.0f);
}

#endif /* CRT_RNG_H */
