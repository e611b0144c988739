#include "scene.h"

#include <stdio.h>
#include <stdlib.h>
#include <string.h>

#include "calc.h"
#include "error.h"
#include "material.h"
#include "object.h"

#include "json.h"

/* Loader context: the file name for messages and the error buffer.  Every
 * helper returns 0 / -1; the first failure's message is kept. */
struct loader {
	const char *file;
	char *err;
};

static int bad(struct loader *ld, const char *fmt, const char *arg)
{
	if (ld->err && !ld->err[0]) {
		char msg[RT_ERRLEN / 2];
		snprintf(msg, sizeof msg, fmt, arg);
		snprintf(ld->err, RT_ERRLEN, "%s in scene [%s].", msg, ld->file);
	}
	return -1;
}

static const char *type_name(enum json_type t)
{
	switch (t) {
	case JSON_NUMBER: return "Number";
	case JSON_STRING: return "String";
	case JSON_ARRAY: return "Array";
	case JSON_OBJECT: return "Object";
	default: return "?";
	}
}

/* member `key` of `parent`, required to have type t */
static const json_value *member(struct loader *ld, const json_value *parent, const char *key, enum json_type t)
{
	const json_value *v = json_get(parent, key);
	if (!json_is(v, t)) {
		if (ld->err && !ld->err[0])
			snprintf(ld->err, RT_ERRLEN, "Expected token [%s] of type [%s] in scene [%s].", key, type_name(t), ld->file);
		return NULL;
	}
	return v;
}

static int floats_from(struct loader *ld, const json_value *arr, float *dst, size_t n)
{
	for (size_t i = 0; i < n; i++) {
		const json_value *e = json_at(arr, i);
		if (!json_is(e, JSON_NUMBER))
			return bad(ld, "Expected token in Array of type [Number]%s", "");
		dst[i] = (float)json_number(e);
	}
	return 0;
}

/* required array member of exactly n numbers */
static int member_floats(struct loader *ld, const json_value *parent, const char *key, float *dst, size_t n)
{
	const json_value *arr = member(ld, parent, key, JSON_ARRAY);
	if (!arr)
		return -1;
	if (json_size(arr) != n) {
		if (ld->err && !ld->err[0])
			snprintf(ld->err, RT_ERRLEN, "Expected token [%s] of length [%zu] in scene [%s].", key, n, ld->file);
		return -1;
	}
	return floats_from(ld, arr, dst, n);
}

static int member_float(struct loader *ld, const json_value *parent, const char *key, float *dst)
{
	const json_value *v = member(ld, parent, key, JSON_NUMBER);
	if (!v)
		return -1;
	*dst = (float)json_number(v);
	return 0;
}

/* ---- camera ------------------------------------------------------------ */

static int load_camera(struct loader *ld, const json_value *j, struct rt_camera *cam)
{
	if (json_size(j) != 5) {
		if (ld->err && !ld->err[0])
			snprintf(ld->err, RT_ERRLEN, "Expected token [Camera] to contain 5 elements.");
		return -1;
	}
	float pos[3], x[3], y[3], fov, focal;
	if (member_floats(ld, j, "position", pos, 3) || member_floats(ld, j, "vector_x", x, 3) ||
		member_floats(ld, j, "vector_y", y, 3) || member_float(ld, j, "fov", &fov) ||
		member_float(ld, j, "focal_length", &focal))
		return -1;
	return rt_camera_init(cam, pos, x, y, fov, focal, ld->err);
}

/* ---- materials --------------------------------------------------------- */

static int load_color_pair(struct loader *ld, const json_value *tex, float colors[2][3])
{
	const json_value *arr = member(ld, tex, "colors", JSON_ARRAY);
	if (!arr)
		return -1;
	if (json_size(arr) != 2)
		return bad(ld, "Expected token [colors] of length [2]%s", "");
	for (size_t i = 0; i < 2; i++) {
		const json_value *c = json_at(arr, i);
		if (!json_is(c, JSON_ARRAY))
			return bad(ld, "Expected token in [colors] of type Array%s", "");
		if (json_size(c) != 3)
			return bad(ld, "Expected token in [colors] of length 3%s", "");
		if (floats_from(ld, c, colors[i], 3))
			return -1;
	}
	return 0;
}

static int load_texture(struct loader *ld, const json_value *tex, crt_material *m)
{
	const json_value *jt = member(ld, tex, "type", JSON_STRING);
	if (!jt)
		return -1;
	const char *type = json_string(jt);
	if (!strcmp(type, "uniform")) {
		m->tex_type = CRT_TEX_UNIFORM;
		return member_floats(ld, tex, "color", m->color[0], 3);
	}
	if (!strcmp(type, "checkerboard")) {
		m->tex_type = CRT_TEX_CHECKERBOARD;
		if (load_color_pair(ld, tex, m->color))
			return -1;
		return member_float(ld, tex, "scale", &m->scale);
	}
	if (!strcmp(type, "brick")) {
		m->tex_type = CRT_TEX_BRICK;
		if (load_color_pair(ld, tex, m->color) || member_float(ld, tex, "scale", &m->scale))
			return -1;
		return member_float(ld, tex, "mortar width", &m->param1);
	}
	if (!strcmp(type, "noisy periodic")) {
		m->tex_type = CRT_TEX_NOISY_PERIODIC;
		if (member_floats(ld, tex, "color", m->color[0], 3) ||
			member_floats(ld, tex, "color gradient", m->color[1], 3) ||
			member_float(ld, tex, "noise feature scale", &m->scale) ||
			member_float(ld, tex, "noise scale", &m->param1) ||
			member_float(ld, tex, "frequency scale", &m->param2))
			return -1;
		const json_value *jf = member(ld, tex, "function", JSON_STRING);
		if (!jf)
			return -1;
		static const char *const names[] = { "sin", "saw", "triangle", "square" };
		for (int f = 0; f < 4; f++)
			if (!strcmp(json_string(jf), names[f])) {
				m->tex_func = f;
				return 0;
			}
		if (ld->err && !ld->err[0])
			snprintf(ld->err, RT_ERRLEN, "Unexpected value [%s] of token [function] in scene [%s].", json_string(jf), ld->file);
		return -1;
	}
	if (ld->err && !ld->err[0])
		snprintf(ld->err, RT_ERRLEN, "Unrecognized token [%s] in texture in scene [%s].", type, ld->file);
	return -1;
}

static int load_material(struct loader *ld, const json_value *j, crt_material *m)
{
	memset(m, 0, sizeof *m);
	const json_value *id = member(ld, j, "id", JSON_NUMBER);
	if (!id)
		return -1;
	m->id = json_int(id);
	if (member_float(ld, j, "shininess", &m->shininess) || member_float(ld, j, "refractive_index", &m->refractive_index))
		return -1;
	const json_value *tex = member(ld, j, "texture", JSON_OBJECT);
	if (!tex)
		return -1;
	if (member_floats(ld, j, "ks", m->ks, 3) || member_floats(ld, j, "ka", m->ka, 3) ||
		member_floats(ld, j, "kr", m->kr, 3) || member_floats(ld, j, "kt", m->kt, 3) ||
		member_floats(ld, j, "ke", m->ke, 3))
		return -1;
	if (load_texture(ld, tex, m))
		return -1;
	rt_material_finalize(m);
	return 0;
}

static int load_materials(struct loader *ld, const json_value *arr, struct rt_scene *sc)
{
	sc->num_materials = json_size(arr);
	if (!sc->num_materials)
		return bad(ld, "Expected token [Materials] to contain nonzero element count%s", "");
	sc->materials = calloc(sc->num_materials, sizeof *sc->materials);
	if (!sc->materials)
		return bad(ld, "Unable to allocate materials%s", "");
	for (size_t i = 0; i < sc->num_materials; i++) {
		const json_value *j = json_at(arr, i);
		if (!json_is(j, JSON_OBJECT))
			return bad(ld, "Expected token in [Materials] of type Object%s", "");
		if (load_material(ld, j, &sc->materials[i]))
			return -1;
	}
	return 0;
}

/* ---- objects ----------------------------------------------------------- */

struct common_params {
	int material; /* index */
	float epsilon;
	unsigned lights;
};

static int load_common(struct loader *ld, const struct rt_scene *sc, const json_value *params, struct common_params *cp)
{
	const json_value *jm = member(ld, params, "material", JSON_NUMBER);
	if (!jm)
		return -1;
	cp->material = rt_material_find(sc->materials, sc->num_materials, json_int(jm));
	if (cp->material < 0) {
		if (ld->err && !ld->err[0])
			snprintf(ld->err, RT_ERRLEN, "Failed to get material id [%d].", json_int(jm));
		return -1;
	}
	const json_value *je = json_get(params, "epsilon");
	const json_value *jl = json_get(params, "lights");
	cp->epsilon = json_is(je, JSON_NUMBER) ? (float)json_number(je) : RT_EPSILON_AUTO;
	cp->lights = json_is(jl, JSON_NUMBER) ? (unsigned)json_int(jl) : 0u;
	return 0;
}

static int append_objects(struct rt_scene *sc, size_t *cap, const crt_object *src, size_t n)
{
	if (sc->num_objects + n > *cap) {
		size_t ncap = *cap ? *cap : 16;
		while (ncap < sc->num_objects + n)
			ncap *= 2;
		crt_object *no = realloc(sc->objects, ncap * sizeof *no);
		if (!no)
			return -1;
		sc->objects = no;
		*cap = ncap;
	}
	memcpy(sc->objects + sc->num_objects, src, n * sizeof *src);
	sc->num_objects += n;
	return 0;
}

static int load_objects(struct loader *ld, const json_value *arr, struct rt_scene *sc)
{
	const size_t n = json_size(arr);
	if (!n)
		return bad(ld, "Expected token [Objects] to contain nonzero element count%s", "");
	size_t cap = 0;
	for (size_t i = 0; i < n; i++) {
		const json_value *j = json_at(arr, i);
		if (!json_is(j, JSON_OBJECT))
			return bad(ld, "Expected token in [Objects] of type Object%s", "");
		const json_value *jt = member(ld, j, "type", JSON_STRING);
		const json_value *params = jt ? member(ld, j, "parameters", JSON_OBJECT) : NULL;
		if (!params)
			return -1;
		struct common_params cp;
		if (load_common(ld, sc, params, &cp))
			return -1;
		const int emits = (sc->materials[cp.material].flags & CRT_MAT_EMITTANT) != 0;
		const char *type = json_string(jt);
		crt_object obj;
		if (!strcmp(type, "Sphere")) {
			float radius, pos[3];
			if (member_float(ld, params, "radius", &radius) || member_floats(ld, params, "position", pos, 3))
				return -1;
			rt_sphere_make(&obj, pos, radius, cp.material, cp.epsilon, cp.lights);
		} else if (!strcmp(type, "Triangle")) {
			float v[3][3];
			if (member_floats(ld, params, "vertex_1", v[0], 3) || member_floats(ld, params, "vertex_2", v[1], 3) ||
				member_floats(ld, params, "vertex_3", v[2], 3))
				return -1;
			rt_triangle_make(&obj, v, cp.material, cp.epsilon, cp.lights);
		} else if (!strcmp(type, "Plane")) {
			float pos[3], normal[3];
			if (member_floats(ld, params, "position", pos, 3) || member_floats(ld, params, "normal", normal, 3))
				return -1;
			if (emits) {
				if (ld->err && !ld->err[0])
					snprintf(ld->err, RT_ERRLEN, "Plane cannot be emittant");
				return -1;
			}
			rt_plane_make(&obj, pos, normal, cp.material, cp.epsilon, cp.lights);
			sc->num_planes++;
		} else if (!strcmp(type, "Mesh")) {
			const json_value *jf = member(ld, params, "filename", JSON_STRING);
			float pos[3], rot[3], scale;
			if (!jf || member_floats(ld, params, "position", pos, 3) || member_floats(ld, params, "rotation", rot, 3) ||
				member_float(ld, params, "scale", &scale))
				return -1;
			/* The reference counts an emissive mesh as an emitter but never
			 * registers its triangles (undefined behaviour). */
			if (emits)
				return bad(ld, "Mesh [%s] cannot be emittant", json_string(jf));
			crt_object *tris;
			size_t count;
			if (rt_stl_load(json_string(jf), pos, rot, scale, cp.material, cp.epsilon, cp.lights, &tris, &count, ld->err))
				return -1;
			const int rc = append_objects(sc, &cap, tris, count);
			free(tris);
			if (rc)
				return bad(ld, "Unable to allocate objects%s", "");
			continue;
		} else {
			return bad(ld, "Unrecognized object type [%s]", type);
		}
		if (emits)
			sc->num_emitters++;
		if (append_objects(sc, &cap, &obj, 1))
			return bad(ld, "Unable to allocate objects%s", "");
	}
	if (!sc->num_emitters)
		return bad(ld, "Expected non-zero number of emittant objects%s", "");
	return 0;
}

/* ---- entry points ------------------------------------------------------ */

void rt_scene_scale(struct rt_scene *sc, float factor)
{
	for (size_t i = 0; i < sc->num_objects; i++)
		rt_object_scale(&sc->objects[i], factor);
	rt_camera_scale(&sc->camera, factor);
}

void rt_scene_free(struct rt_scene *sc)
{
	free(sc->materials);
	free(sc->objects);
	memset(sc, 0, sizeof *sc);
}

void rt_scene_desc(const struct rt_scene *sc, crt_scene_desc *desc)
{
	memset(desc, 0, sizeof *desc);
	desc->abi_version = CRT_ABI_VERSION;
	desc->num_objects = (uint32_t)sc->num_objects;
	desc->objects = sc->objects;
	desc->num_materials = (uint32_t)sc->num_materials;
	desc->materials = sc->materials;
	v3_copy(desc->ambient, sc->ambient);
	desc->build = sc->build;
}

static char *slurp(const char *filename, char *err)
{
	FILE *f = fopen(filename, "rb");
	if (!f) {
		snprintf(err, RT_ERRLEN, "Unable to open scene file [%s].", filename);
		return NULL;
	}
	fseek(f, 0, SEEK_END);
	const long len = ftell(f);
	fseek(f, 0, SEEK_SET);
	char *text = malloc(len >= 0 ? (size_t)len + 1 : 1);
	if (!text || len < 0 || fread(text, 1, (size_t)len, f) != (size_t)len) {
		snprintf(err, RT_ERRLEN, "Failed to read scene file [%s].", filename);
		free(text);
		fclose(f);
		return NULL;
	}
	fclose(f);
	text[len] = '\0';
	return text;
}

int rt_scene_load(const char *filename, const char *rescale, struct rt_scene *sc, char *err)
{
	char local_err[RT_ERRLEN];
	if (!err)
		err = local_err;
	err[0] = '\0';
	memset(sc, 0, sizeof *sc);

	char *text = slurp(filename, err);
	if (!text)
		return -1;
	char parse_err[128];
	json_value *root = json_parse(text, parse_err, sizeof parse_err);
	free(text);
	if (!root)
		rt_fail(err, "Failed to parse scene [%s]: %s.", filename, parse_err);
	if (!json_is(root, JSON_OBJECT)) {
		json_free(root);
		rt_fail(err, "Expected parent token of type Object in scene [%s].", filename);
	}

	struct loader ld = { filename, err };
	const json_value *materials = member(&ld, root, "Materials", JSON_ARRAY);
	const json_value *objects = materials ? member(&ld, root, "Objects", JSON_ARRAY) : NULL;
	const json_value *camera = objects ? member(&ld, root, "Camera", JSON_OBJECT) : NULL;
	int rc = camera ? 0 : -1;
	if (!rc)
		rc = load_camera(&ld, camera, &sc->camera);
	if (!rc)
		rc = load_materials(&ld, materials, sc);
	if (!rc)
		rc = load_objects(&ld, objects, sc);
	if (!rc) {
		const json_value *amb = json_get(root, "AmbientLight");
		if (json_is(amb, JSON_ARRAY) && json_size(amb) == 3)
			rc = floats_from(&ld, amb, sc->ambient, 3);
	}
	json_free(root);
	if (rc) {
		rt_scene_free(sc);
		return -1;
	}

	if (rescale) {
		if (!strcmp(rescale, "norm")) {
			float lo[3], hi[3];
			rt_objects_extents(sc->objects, sc->num_objects, lo, hi);
			float range = hi[0] - lo[0];
			if (range < hi[1] - lo[1])
				range = hi[1] - lo[1];
			if (range < hi[2] - lo[2])
				range = hi[2] - lo[2];
			rt_scene_scale(sc, 1.f / range);
		} else {
			rt_scene_scale(sc, (float)atof(rescale));
		}
	}
	return 0;
}
