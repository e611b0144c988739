/* rest_box_emu -- the device's "does this shadow ray miss the box of the other objects" test
 * (c-raytracer_b200/csrc/crt_device.cuh: ray_misses_box, safe_idir / make_wide_ray; the box as
 * c-raytracer_b200/csrc/crt_api.cu pads it) restated on the CPU in the same FP32 operations (compile with
 * -ffp-contract=off: the device code has no contraction opportunity in it either), fuzzed for FALSE CULLS:
 *     rest_box_emu <cases> <seed>
 * A case = a float ray (origin, unit direction as the kernel forms it from a light point), a distance t
 * along it where an occluder is met, a light distance dist > t, and a box that contains the point
 * origin + t * direction (computed in double).  Such a ray meets the box within [0, dist], so the test
 * must NOT report a miss -- a miss would skip the tree walk and lose the occluder.
 * Kinds: random; zero direction components; origins 10..10^6 box sizes away; flat and tiny boxes; the
 * point on a face / edge / corner of the box; origin inside the box; dist barely beyond t.
 * Also counts how selective the test is (rays aimed beside the box that it does cull).
 * Exits 1 on the first false cull. */
#include <cfloat>
#include <cmath>
#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <random>

/* ---- restatement of the device code ------------------------------------------------------- */
static float safe_idir(float d, float maxd)
{
	const float a = std::fabs(d) > 1e-18f ? d : std::copysign(1e-18f, d);
	return maxd / a;
}

static bool ray_misses_box(const float lo[3], const float hi[3], const float o[3], const float d[3], float dist)
{
	const float maxd = std::fmax(std::fmax(std::fabs(d[0]), std::fabs(d[1])), std::fabs(d[2]));
	const float inx = safe_idir(d[0], maxd), iny = safe_idir(d[1], maxd), inz = safe_idir(d[2], maxd);
	if (lo[0] > hi[0])
		return true;
	const float ax = (lo[0] - o[0]) * inx, bx = (hi[0] - o[0]) * inx;
	const float ay = (lo[1] - o[1]) * iny, by = (hi[1] - o[1]) * iny;
	const float az = (lo[2] - o[2]) * inz, bz = (hi[2] - o[2]) * inz;
	const float tn = fmaxf(fmaxf(fminf(ax, bx), fminf(ay, by)), fminf(az, bz));
	const float tf = fminf(fminf(fmaxf(ax, bx), fmaxf(ay, by)), fmaxf(az, bz));
	const float lim = dist * maxd;
	const float slack = fmaf(fabsf(tn) + fabsf(tf), 1e-5f, 1e-30f);
	return tn - tf > slack || tf < -slack || tn > fmaf(fabsf(lim), 1e-5f, lim) + slack;
}

/* the host's padding of the box (crt_api.cu, DEmitter.rest_lo / rest_hi) */
static void pad_box(float lo[3], float hi[3])
{
	for (int a = 0; a < 3; a++) {
		const float l = lo[a], u = hi[a];
		const float pad = 1e-5f * std::fmax(1e-3f, std::fmax(std::fabs(l), std::fabs(u))) + 1e-5f * (u - l);
		lo[a] = l - pad;
		hi[a] = u + pad;
	}
}

/* ---- the fuzzer ----------------------------------------------------------------------------- */
int main(int argc, char **argv)
{
	const long cases = argc > 1 ? atol(argv[1]) : 2000000;
	std::mt19937_64 rng(argc > 2 ? strtoull(argv[2], nullptr, 10) : 1);
	std::uniform_real_distribution<double> U(0.0, 1.0);
	auto sym = [&]() { return 2.0 * U(rng) - 1.0; };
	long tested = 0, beside = 0, beside_culled = 0;
	for (long c = 0; c < cases; c++) {
		const int kind = (int)(rng() % 8);
		/* scene scale and box size */
		const double scale = std::pow(10.0, 3.0 * sym());            /* 1e-3 .. 1e3 */
		double size[3];
		for (int a = 0; a < 3; a++)
			size[a] = scale * std::pow(10.0, -2.0 * U(rng));          /* up to 100 : 1 aspect */
		if (kind == 3)
			size[rng() % 3] = 0.0;                                    /* flat box */
		/* origin: inside, near, or far */
		const double far = kind == 2 ? std::pow(10.0, 1.0 + 5.0 * U(rng)) : (kind == 5 ? 0.3 : 1.0 + 9.0 * U(rng));
		float o[3], centre[3];
		for (int a = 0; a < 3; a++)
			centre[a] = (float)(scale * 10.0 * sym());
		double dir[3] = { sym(), sym(), sym() };
		if (kind == 1) { /* axis-parallel or in an axis plane */
			dir[rng() % 3] = 0.0;
			if (rng() & 1)
				dir[rng() % 3] = 0.0;
			if (dir[0] == 0.0 && dir[1] == 0.0 && dir[2] == 0.0)
				dir[rng() % 3] = 1.0;
		}
		const double dn = std::sqrt(dir[0] * dir[0] + dir[1] * dir[1] + dir[2] * dir[2]);
		if (dn < 1e-3)
			continue;
		const double reach = far * scale;
		for (int a = 0; a < 3; a++)
			o[a] = (float)(centre[a] - dir[a] / dn * reach);
		/* the kernel's direction: v = L - P, d2, dist = sqrtf(d2), inv = 1 / dist, d = v * inv (crt_kernels.cu) */
		float L[3], v[3], d[3];
		for (int a = 0; a < 3; a++) {
			L[a] = (float)(o[a] + dir[a] / dn * reach * (1.5 + U(rng)));
			v[a] = L[a] - o[a];
		}
		const float d2 = v[0] * v[0] + v[1] * v[1] + v[2] * v[2];
		const float dist = std::sqrt(d2);
		if (!(dist > 0.f) || !std::isfinite(dist))
			continue;
		const float inv = 1.f / dist;
		for (int a = 0; a < 3; a++)
			d[a] = v[a] * inv;
		/* where the occluder is met: t in (0, dist) */
		double t = (double)dist * (kind == 6 ? 1.0 - 1e-6 * (1.0 + U(rng)) : 0.05 + 0.9 * U(rng));
		double X[3];
		for (int a = 0; a < 3; a++)
			X[a] = (double)o[a] + t * (double)d[a];
		/* a box that contains X: X anywhere inside, or on a face / edge / corner (kind 4) */
		float lo[3], hi[3];
		for (int a = 0; a < 3; a++) {
			double f = U(rng);
			if (kind == 4 && (rng() & 1))
				f = (rng() & 1) ? 0.0 : 1.0;
			const double l = X[a] - f * size[a], u = X[a] + (1.0 - f) * size[a];
			lo[a] = std::nextafter((float)l, -FLT_MAX); /* outward to float: the box of real primitives contains them */
			hi[a] = std::nextafter((float)u, FLT_MAX);
			if ((double)lo[a] > l) lo[a] = std::nextafter(lo[a], -FLT_MAX);
			if ((double)hi[a] < u) hi[a] = std::nextafter(hi[a], FLT_MAX);
		}
		float plo[3] = { lo[0], lo[1], lo[2] }, phi[3] = { hi[0], hi[1], hi[2] };
		pad_box(plo, phi);
		tested++;
		if (ray_misses_box(plo, phi, o, d, dist)) {
			printf("FALSE CULL (case %ld kind %d): o %.9g %.9g %.9g d %.9g %.9g %.9g dist %.9g t %.17g box [%.9g %.9g %.9g] [%.9g %.9g %.9g]\n",
				c, kind, o[0], o[1], o[2], d[0], d[1], d[2], dist, t, lo[0], lo[1], lo[2], hi[0], hi[1], hi[2]);
			return 1;
		}
		/* selectivity: the same box moved aside by three of its sizes across the ray */
		{
			int m = 0; /* an axis the ray is not mainly along */
			for (int a = 1; a < 3; a++)
				if (std::fabs(d[a]) < std::fabs(d[m]))
					m = a;
			const float shift = (float)(3.0 * size[m] + 0.01 * scale);
			float slo[3] = { lo[0], lo[1], lo[2] }, shi[3] = { hi[0], hi[1], hi[2] };
			slo[m] += shift + (hi[m] - lo[m]);
			shi[m] += shift + (hi[m] - lo[m]);
			pad_box(slo, shi);
			beside++;
			if (ray_misses_box(slo, shi, o, d, dist))
				beside_culled++;
		}
	}
	/* a NaN ray is never culled (it walks the tree as before), an empty box always is */
	{
		const float lo[3] = { -1.f, -1.f, -1.f }, hi[3] = { 1.f, 1.f, 1.f }, o[3] = { 5.f, 0.f, 0.f };
		const float nan = std::nanf("");
		const float dn[3] = { nan, nan, nan };
		if (ray_misses_box(lo, hi, o, dn, nan) || ray_misses_box(lo, hi, o, dn, 3.f)) {
			printf("NaN ray culled\n");
			return 1;
		}
		const float elo[3] = { 1.f, 0.f, 0.f }, ehi[3] = { -1.f, 0.f, 0.f }, dx[3] = { -1.f, 0.f, 0.f };
		if (!ray_misses_box(elo, ehi, o, dx, 10.f)) {
			printf("empty box not culled\n");
			return 1;
		}
		/* box behind the origin, and box beyond the light */
		const float away[3] = { 1.f, 0.f, 0.f };
		if (!ray_misses_box(lo, hi, o, away, 10.f) || !ray_misses_box(lo, hi, o, dx, 3.9f) || ray_misses_box(lo, hi, o, dx, 4.1f)) {
			printf("behind / beyond cases wrong\n");
			return 1;
		}
	}
	printf("%ld rays that meet their box: 0 false culls; %ld boxes beside the ray: %ld culled (%.1f %%)\n", tested, beside, beside_culled,
		100.0 * (double)beside_culled / (double)(beside ? beside : 1));
	return 0;
}
