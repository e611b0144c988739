/*
 * json.c -- recursive-descent JSON reader.  See json.h.
 */
#include "json.h"

#include <limits.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

struct json_value {
	enum json_type type;
	double number;
	char *string;         /* JSON_STRING */
	size_t count;         /* elements / members */
	json_value **items;   /* JSON_ARRAY / JSON_OBJECT values */
	char **keys;          /* JSON_OBJECT */
};

struct parser {
	const char *text;
	const char *p;
	char *err;
	size_t errlen;
	int depth;
};

#define JSON_MAX_DEPTH 512

static json_value *parse_value(struct parser *ps);

static void fail(struct parser *ps, const char *what)
{
	if (ps->err && ps->errlen && !ps->err[0])
		snprintf(ps->err, ps->errlen, "%s at byte %zu", what, (size_t)(ps->p - ps->text));
}

static void skip_ws(struct parser *ps)
{
	while (*ps->p == ' ' || *ps->p == '\t' || *ps->p == '\n' || *ps->p == '\r')
		ps->p++;
}

static json_value *new_value(enum json_type t)
{
	json_value *v = calloc(1, sizeof *v);
	if (v)
		v->type = t;
	return v;
}

static int hex4(const char *s, unsigned *out)
{
	unsigned v = 0;
	for (int i = 0; i < 4; i++) {
		char c = s[i];
		v <<= 4;
		if (c >= '0' && c <= '9')
			v |= (unsigned)(c - '0');
		else if (c >= 'a' && c <= 'f')
			v |= (unsigned)(c - 'a' + 10);
		else if (c >= 'A' && c <= 'F')
			v |= (unsigned)(c - 'A' + 10);
		else
			return 0;
	}
	*out = v;
	return 1;
}

static size_t utf8_put(char *dst, unsigned cp)
{
	if (cp < 0x80) {
		dst[0] = (char)cp;
		return 1;
	}
	if (cp < 0x800) {
		dst[0] = (char)(0xC0 | (cp >> 6));
		dst[1] = (char)(0x80 | (cp & 0x3F));
		return 2;
	}
	if (cp < 0x10000) {
		dst[0] = (char)(0xE0 | (cp >> 12));
		dst[1] = (char)(0x80 | ((cp >> 6) & 0x3F));
		dst[2] = (char)(0x80 | (cp & 0x3F));
		return 3;
	}
	dst[0] = (char)(0xF0 | (cp >> 18));
	dst[1] = (char)(0x80 | ((cp >> 12) & 0x3F));
	dst[2] = (char)(0x80 | ((cp >> 6) & 0x3F));
	dst[3] = (char)(0x80 | (cp & 0x3F));
	return 4;
}

/* ps->p points at the opening quote; returns a malloc'ed decoded string */
static char *parse_string_raw(struct parser *ps)
{
	const char *s = ps->p + 1;
	const char *e = s;
	while (*e && *e != '"') {
		if ((unsigned char)*e < 0x20) {
			ps->p = e;
			fail(ps, "control character in string");
			return NULL;
		}
		if (*e == '\\' && e[1])
			e++;
		e++;
	}
	if (*e != '"') {
		ps->p = e;
		fail(ps, "unterminated string");
		return NULL;
	}
	char *out = malloc((size_t)(e - s) + 1);
	if (!out) {
		fail(ps, "out of memory");
		return NULL;
	}
	size_t n = 0;
	while (s < e) {
		if (*s != '\\') {
			out[n++] = *s++;
			continue;
		}
		s++;
		switch (*s) {
		case '"': out[n++] = '"'; break;
		case '\\': out[n++] = '\\'; break;
		case '/': out[n++] = '/'; break;
		case 'b': out[n++] = '\b'; break;
		case 'f': out[n++] = '\f'; break;
		case 'n': out[n++] = '\n'; break;
		case 'r': out[n++] = '\r'; break;
		case 't': out[n++] = '\t'; break;
		case 'u': {
			unsigned cp;
			if (e - s < 5 || !hex4(s + 1, &cp))
				goto bad_escape;
			s += 4;
			if (cp >= 0xD800 && cp <= 0xDBFF) { /* surrogate pair */
				unsigned lo;
				if (e - s < 7 || s[1] != '\\' || s[2] != 'u' || !hex4(s + 3, &lo) || lo < 0xDC00 || lo > 0xDFFF)
					goto bad_escape;
				cp = 0x10000 + (((cp - 0xD800) << 10) | (lo - 0xDC00));
				s += 6;
			} else if (cp >= 0xDC00 && cp <= 0xDFFF) {
				goto bad_escape;
			}
			n += utf8_put(out + n, cp);
		} break;
		default:
			goto bad_escape;
		}
		s++;
	}
	out[n] = '\0';
	ps->p = e + 1;
	return out;
bad_escape:
	ps->p = s;
	fail(ps, "invalid escape sequence");
	free(out);
	return NULL;
}

static json_value *parse_number(struct parser *ps)
{
	const char *s = ps->p;
	const char *q = s;
	if (*q == '-')
		q++;
	if (*q == '0') {
		q++;
	} else if (*q >= '1' && *q <= '9') {
		while (*q >= '0' && *q <= '9')
			q++;
	} else {
		fail(ps, "invalid number");
		return NULL;
	}
	if (*q == '.') {
		q++;
		if (!(*q >= '0' && *q <= '9')) {
			ps->p = q;
			fail(ps, "invalid number");
			return NULL;
		}
		while (*q >= '0' && *q <= '9')
			q++;
	}
	if (*q == 'e' || *q == 'E') {
		q++;
		if (*q == '+' || *q == '-')
			q++;
		if (!(*q >= '0' && *q <= '9')) {
			ps->p = q;
			fail(ps, "invalid number");
			return NULL;
		}
		while (*q >= '0' && *q <= '9')
			q++;
	}
	char tmp[64];
	size_t len = (size_t)(q - s);
	double d;
	if (len < sizeof tmp) {
		memcpy(tmp, s, len);
		tmp[len] = '\0';
		d = strtod(tmp, NULL);
	} else {
		char *big = malloc(len + 1);
		if (!big) {
			fail(ps, "out of memory");
			return NULL;
		}
		memcpy(big, s, len);
		big[len] = '\0';
		d = strtod(big, NULL);
		free(big);
	}
	json_value *v = new_value(JSON_NUMBER);
	if (!v) {
		fail(ps, "out of memory");
		return NULL;
	}
	v->number = d;
	ps->p = q;
	return v;
}

static int push_item(json_value *c, size_t *cap, json_value *item, char *key)
{
	if (c->count == *cap) {
		size_t ncap = *cap ? *cap * 2 : 8;
		json_value **ni = realloc(c->items, ncap * sizeof *ni);
		if (!ni)
			return 0;
		c->items = ni;
		if (c->type == JSON_OBJECT) {
			char **nk = realloc(c->keys, ncap * sizeof *nk);
			if (!nk)
				return 0;
			c->keys = nk;
		}
		*cap = ncap;
	}
	c->items[c->count] = item;
	if (c->type == JSON_OBJECT)
		c->keys[c->count] = key;
	c->count++;
	return 1;
}

static json_value *parse_container(struct parser *ps, enum json_type type)
{
	const char close = type == JSON_ARRAY ? ']' : '}';
	json_value *c = new_value(type);
	size_t cap = 0;
	if (!c) {
		fail(ps, "out of memory");
		return NULL;
	}
	if (++ps->depth > JSON_MAX_DEPTH) {
		fail(ps, "nesting too deep");
		goto error;
	}
	ps->p++; /* opening bracket */
	skip_ws(ps);
	if (*ps->p == close) {
		ps->p++;
		ps->depth--;
		return c;
	}
	for (;;) {
		char *key = NULL;
		skip_ws(ps);
		if (type == JSON_OBJECT) {
			if (*ps->p != '"') {
				fail(ps, "expected member name");
				goto error;
			}
			key = parse_string_raw(ps);
			if (!key)
				goto error;
			skip_ws(ps);
			if (*ps->p != ':') {
				free(key);
				fail(ps, "expected ':'");
				goto error;
			}
			ps->p++;
		}
		json_value *item = parse_value(ps);
		if (!item) {
			free(key);
			goto error;
		}
		if (!push_item(c, &cap, item, key)) {
			free(key);
			json_free(item);
			fail(ps, "out of memory");
			goto error;
		}
		skip_ws(ps);
		if (*ps->p == ',') {
			ps->p++;
			continue;
		}
		if (*ps->p == close) {
			ps->p++;
			break;
		}
		fail(ps, "expected ',' or closing bracket");
		goto error;
	}
	ps->depth--;
	return c;
error:
	json_free(c);
	return NULL;
}

static json_value *parse_literal(struct parser *ps, const char *word, enum json_type t, double num)
{
	size_t n = strlen(word);
	if (strncmp(ps->p, word, n)) {
		fail(ps, "unexpected token");
		return NULL;
	}
	json_value *v = new_value(t);
	if (!v) {
		fail(ps, "out of memory");
		return NULL;
	}
	v->number = num;
	ps->p += n;
	return v;
}

static json_value *parse_value(struct parser *ps)
{
	skip_ws(ps);
	switch (*ps->p) {
	case '{':
		return parse_container(ps, JSON_OBJECT);
	case '[':
		return parse_container(ps, JSON_ARRAY);
	case '"': {
		char *s = parse_string_raw(ps);
		if (!s)
			return NULL;
		json_value *v = new_value(JSON_STRING);
		if (!v) {
			free(s);
			fail(ps, "out of memory");
			return NULL;
		}
		v->string = s;
		return v;
	}
	case 't':
		return parse_literal(ps, "true", JSON_BOOL, 1.0);
	case 'f':
		return parse_literal(ps, "false", JSON_BOOL, 0.0);
	case 'n':
		return parse_literal(ps, "null", JSON_NULL, 0.0);
	case '\0':
		fail(ps, "unexpected end of input");
		return NULL;
	default:
		return parse_number(ps);
	}
}

json_value *json_parse(const char *text, char *err, size_t errlen)
{
	struct parser ps = { text, text, err, errlen, 0 };
	if (err && errlen)
		err[0] = '\0';
	if (!text) {
		fail(&ps, "no input");
		return NULL;
	}
	/* tolerate a UTF-8 byte-order mark, as cJSON does */
	if (!strncmp(ps.p, "\xEF\xBB\xBF", 3))
		ps.p += 3;
	json_value *v = parse_value(&ps);
	if (!v)
		return NULL;
	skip_ws(&ps);
	if (*ps.p) {
		fail(&ps, "trailing characters after document");
		json_free(v);
		return NULL;
	}
	return v;
}

void json_free(json_value *v)
{
	if (!v)
		return;
	for (size_t i = 0; i < v->count; i++) {
		json_free(v->items[i]);
		if (v->keys)
			free(v->keys[i]);
	}
	free(v->items);
	free(v->keys);
	free(v->string);
	free(v);
}

enum json_type json_typeof(const json_value *v)
{
	return v ? v->type : JSON_NULL;
}

int json_is(const json_value *v, enum json_type t)
{
	return v && v->type == t;
}

size_t json_size(const json_value *v)
{
	return (v && (v->type == JSON_ARRAY || v->type == JSON_OBJECT)) ? v->count : 0;
}

const json_value *json_at(const json_value *v, size_t i)
{
	return i < json_size(v) ? v->items[i] : NULL;
}

const char *json_key_at(const json_value *v, size_t i)
{
	return (v && v->type == JSON_OBJECT && i < v->count) ? v->keys[i] : NULL;
}

const json_value *json_get(const json_value *v, const char *key)
{
	if (!v || v->type != JSON_OBJECT || !key)
		return NULL;
	for (size_t i = 0; i < v->count; i++)
		if (!strcmp(v->keys[i], key))
			return v->items[i];
	return NULL;
}

double json_number(const json_value *v)
{
	return (v && v->type == JSON_NUMBER) ? v->number : 0.0;
}

int json_int(const json_value *v)
{
	double d = json_number(v);
	if (d >= (double)INT_MAX)
		return INT_MAX;
	if (d <= (double)INT_MIN)
		return INT_MIN;
	return (int)d;
}

const char *json_string(const json_value *v)
{
	return (v && v->type == JSON_STRING) ? v->string : NULL;
}

int json_bool(const json_value *v)
{
	return v && v->type == JSON_BOOL && v->number != 0.0;
}
