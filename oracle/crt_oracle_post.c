/*
 * crt_oracle_post.c -- TEST INFRASTRUCTURE.  CPU restatement of the reference's
 * post-processing operations (src/postprocess/postproc.c) behind the same
 * options struct as crt_postprocess() of include/crt.h.  Only tests/,
 * __graft_entry__.smoke() and the benchmarks' cpu_baseline legs may call it;
 * the product path (c-raytracer_b200/csrc/crt_post.cu) never does.
 *
 * Parity pinned: checked bit for bit against the unmodified reference
 * (oracle/_ref/ref_post, built from postproc.c + image.c with IEEE flags) in
 * tests/test_oracle_post.py, on frames rendered by the reference engine.
 *
 * Compiled -O2 -ffp-contract=off: every product and sum rounds once, as the
 * reference's scalar C does without -ffast-math.
 */
#include <float.h>
#include <math.h>
#include <stdint.h>
#include <stdlib.h>
#include <string.h>

#include "crt.h"

/* brighten(), postproc.c:94-102 */
static void post_brighten(float *raster, size_t pixels, float factor)
{
	for (size_t i = 0; i < 3 * pixels; i++)
		raster[i] = raster[i] * factor;
}

/* clamp(), calc.c:164-168 */
static float clampf(float v, float lo, float hi)
{
	const float r = v < lo ? lo : v;
	return r > hi ? hi : r;
}

/* mist(), postproc.c:162-188 */
static void post_mist(float *raster, const float *z, size_t pixels, float start, float depth, uint32_t falloff, const float color[3])
{
	const float inv_depth = 1.f / depth;
	for (size_t i = 0; i < pixels; i++) {
		const float factor = clampf((z[i] - start) * inv_depth, 0.f, 1.f);
		float opacity = factor;
		if (falloff == CRT_FALLOFF_QUADRATIC)
			opacity = factor * factor;
		else if (falloff == CRT_FALLOFF_INV_QUADRATIC)
			opacity = sqrtf(factor);
		const float keep = 1.f - opacity;
		for (int c = 0; c < 3; c++) {
			const float kept = raster[3 * i + c] * keep;
			const float added = color[c] * opacity;
			raster[3 * i + c] = kept + added;
		}
	}
}

/* depth_of_field(), postproc.c:104-160: every source pixel, in index order,
 * spreads colour * alpha over its circle of confusion (column by column, the
 * column's half height is (int)sqrtf(r^2 - x^2)), clipped to the frame, onto
 * the pixels that are not in front of it; then the sums are normalised.
 * Returns the largest radius, or -1 when a radius does not fit the reference's
 * int arithmetic. */
static int post_dof(float *raster, const float *z, int w, int h, float scale, float bias)
{
	const size_t pixels = (size_t)w * h;
	float *sum = calloc(pixels, 3 * sizeof(float));
	float *alpha_sum = calloc(pixels, sizeof(float));
	int rmax = 0;
	for (size_t i = 0; i < pixels; i++) {
		const float depth = z[i];
		const float coc = fabsf(depth * scale + bias);
		const float half = coc * .5f;
		if (!(half < 32767.f)) {
			free(sum);
			free(alpha_sum);
			return -1;
		}
		const int r = (int)half, rr = r * r;
		if (r > rmax)
			rmax = r;
		const float inv = 1.f / rr;
		const float alpha = inv < 1.f ? inv : 1.f;
		const float pr = raster[3 * i] * alpha, pg = raster[3 * i + 1] * alpha, pb = raster[3 * i + 2] * alpha;
		const int cx = (int)(i % (size_t)w), cy = (int)(i / (size_t)w);
		const int x0 = cx - r < 0 ? -cx : -r, x1 = cx + r >= w ? w - 1 - cx : r;
		for (int x = x0; x <= x1; x++) {
			const int hh = (int)sqrtf((float)(rr - x * x));
			const int y0 = cy - hh < 0 ? 0 : cy - hh, y1 = cy + hh >= h ? h - 1 : cy + hh;
			for (int y = y0; y <= y1; y++) {
				const size_t t = (size_t)y * w + (size_t)(cx + x);
				if (depth <= z[t]) {
					sum[3 * t] += pr;
					sum[3 * t + 1] += pg;
					sum[3 * t + 2] += pb;
					alpha_sum[t] += alpha;
				}
			}
		}
	}
	for (size_t i = 0; i < pixels; i++) {
		const float inv = 1.f / alpha_sum[i];
		raster[3 * i] = sum[3 * i] * inv;
		raster[3 * i + 1] = sum[3 * i + 1] * inv;
		raster[3 * i + 2] = sum[3 * i + 2] * inv;
	}
	free(sum);
	free(alpha_sum);
	return rmax;
}

/* postprocess(), postproc.c:36-92: brighten, then one of the two depth-of-field
 * forms, then mist */
int oracle_postprocess(const crt_post_opts *o, uint32_t width, uint32_t height, float *raster, const float *z, crt_post_stats *stats)
{
	if (!o || !raster || !z || !width || !height)
		return CRT_ERR_INVALID;
	const size_t pixels = (size_t)width * height;
	crt_post_stats st;
	memset(&st, 0, sizeof st);
	if (o->ops & CRT_POST_BRIGHTEN)
		post_brighten(raster, pixels, o->brighten);
	if (o->ops & (CRT_POST_DOF | CRT_POST_DOF_CAMERA)) {
		float scale = o->dof_scale, bias = o->dof_bias;
		if (!(o->ops & CRT_POST_DOF)) {
			float z_min = FLT_MAX, z_max = FLT_MIN; /* postproc.c:53 */
			for (size_t i = 0; i < pixels; i++) {
				if (z[i] < z_min)
					z_min = z[i];
				if (z[i] > z_max)
					z_max = z[i];
			}
			const float aperture = o->dof_aperture, focal_length = o->dof_focal_length, plane_in_focus = o->dof_plane_in_focus;
			scale = (aperture * focal_length * plane_in_focus * (z_max - z_min)) / ((plane_in_focus - focal_length) * z_min * z_max);
			bias = (aperture * focal_length * (z_min - plane_in_focus)) / ((plane_in_focus * focal_length) * z_min);
			if (!isfinite(scale) || !isfinite(bias))
				return CRT_ERR_INVALID;
		}
		st.dof_scale = scale;
		st.dof_bias = bias;
		const int rmax = post_dof(raster, z, (int)width, (int)height, scale, bias);
		if (rmax < 0)
			return CRT_ERR_UNSUPPORTED;
		st.dof_max_radius = (uint32_t)rmax;
	}
	if (o->ops & CRT_POST_MIST) {
		if (o->mist_falloff > CRT_FALLOFF_INV_QUADRATIC)
			return CRT_ERR_INVALID;
		post_mist(raster, z, pixels, o->mist_start, o->mist_depth, o->mist_falloff, o->mist_color);
	}
	if (stats)
		*stats = st;
	return CRT_OK;
}
