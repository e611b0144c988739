/*
 * json.h -- small strict JSON (RFC 8259) reader used by the scene loader.
 *
 * The reference reads scenes through a vendored cJSON 1.7.14
 * (reference lib/cJSON, used from src/raytracer/scene.c:33-66, :87).  This is
 * an independent implementation of the subset of behaviour scene.c relies on:
 * case-sensitive member lookup returning the first match, member/element
 * counts, numbers as double with a saturating int view, strings with escape
 * decoding.  It is a tree of immutable nodes freed with json_free().
 */
#ifndef RT_JSON_H
#define RT_JSON_H

#include <stddef.h>

enum json_type {
	JSON_NULL = 0,
	JSON_BOOL,
	JSON_NUMBER,
	JSON_STRING,
	JSON_ARRAY,
	JSON_OBJECT,
};

typedef struct json_value json_value;

/* Parse a NUL-terminated text.  Returns NULL on error and writes a message
 * (with byte offset) into err if errlen > 0. */
json_value *json_parse(const char *text, char *err, size_t errlen);
void json_free(json_value *v);

enum json_type json_typeof(const json_value *v); /* JSON_NULL for v == NULL */
int json_is(const json_value *v, enum json_type t);

/* number of elements (array) or members (object); 0 otherwise */
size_t json_size(const json_value *v);
/* i-th element / member value, NULL if out of range */
const json_value *json_at(const json_value *v, size_t i);
/* key of the i-th member of an object, NULL otherwise */
const char *json_key_at(const json_value *v, size_t i);
/* first member named key (case-sensitive), NULL if absent or v not an object */
const json_value *json_get(const json_value *v, const char *key);

double json_number(const json_value *v);   /* 0 if not a number */
int json_int(const json_value *v);         /* saturating (int) view of the number */
const char *json_string(const json_value *v); /* NULL if not a string */
int json_bool(const json_value *v);

#endif /* RT_JSON_H */
