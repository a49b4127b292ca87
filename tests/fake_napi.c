/*
 * Test double for the Node-API functions napi/apm_b200_addon.c uses, plus a small "script" API
 * (fk_*) so a ctypes test can build JavaScript-like values, load the addon's module and call its
 * methods in-process.  Values are never freed (tests are short-lived).
 */
#include <stdbool.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

#include "../napi/node_api_min.h"

typedef enum { FK_UNDEF, FK_NUMBER, FK_STRING, FK_OBJECT, FK_ARRAY, FK_TYPED, FK_BUFFER, FK_FUNCTION, FK_CLASS } fk_kind;

typedef struct fk_prop { char* name; struct fk_value* v; struct fk_prop* next; } fk_prop;
typedef struct fk_value {
    fk_kind kind;
    double num;
    char* str;
    fk_prop* props;                 /* object properties / class methods */
    struct fk_value** elems;        /* array */
    uint32_t nelems;
    napi_typedarray_type ttype;     /* typed array */
    void* data;
    size_t len;                     /* typed array elements / buffer bytes */
    napi_callback cb;               /* function / constructor */
    void* wrapped;
    struct fk_value* cls;           /* instance -> its class */
} fk_value;

struct napi_callback_info__ { size_t argc; fk_value** argv; fk_value* self; };
static char g_exception[1024];
static int g_env_tag;

static fk_value* mk(fk_kind k) { fk_value* v = (fk_value*)calloc(1, sizeof *v); v->kind = k; return v; }
static fk_value* prop_get(fk_value* o, const char* name) {
    for (fk_prop* p = o->props; p; p = p->next) if (!strcmp(p->name, name)) return p->v;
    return NULL;
}
static void prop_set(fk_value* o, const char* name, fk_value* v) {
    for (fk_prop* p = o->props; p; p = p->next) if (!strcmp(p->name, name)) { p->v = v; return; }
    fk_prop* p = (fk_prop*)calloc(1, sizeof *p);
    p->name = strdup(name); p->v = v; p->next = o->props; o->props = p;
}
#define V(x) ((fk_value*)(x))
#define N(x) ((napi_value)(x))

/* ---------------- script API for the test driver ---------------- */
void* fk_env(void) { return &g_env_tag; }
void* fk_number(double d) { fk_value* v = mk(FK_NUMBER); v->num = d; return v; }
void* fk_string(const char* s) { fk_value* v = mk(FK_STRING); v->str = strdup(s); return v; }
void* fk_object(void) { return mk(FK_OBJECT); }
void fk_set(void* o, const char* name, void* v) { prop_set(V(o), name, V(v)); }
void* fk_get(void* o, const char* name) { return prop_get(V(o), name); }
void* fk_array(const double* vals, uint32_t n) {
    fk_value* a = mk(FK_ARRAY);
    a->elems = (fk_value**)calloc(n ? n : 1, sizeof(fk_value*));
    a->nelems = n;
    for (uint32_t i = 0; i < n; ++i) a->elems[i] = V(fk_number(vals[i]));
    return a;
}
void* fk_array_of(void** vals, uint32_t n) {
    fk_value* a = mk(FK_ARRAY);
    a->elems = (fk_value**)calloc(n ? n : 1, sizeof(fk_value*));
    a->nelems = n;
    for (uint32_t i = 0; i < n; ++i) a->elems[i] = V(vals[i]);
    return a;
}
void* fk_typedarray(int type, void* data, size_t n) {
    fk_value* t = mk(FK_TYPED); t->ttype = (napi_typedarray_type)type; t->data = data; t->len = n; return t;
}
int fk_kind_of(void* v) { return v ? (int)V(v)->kind : -1; }
double fk_number_value(void* v) { return V(v)->num; }
const char* fk_string_value(void* v) { return V(v)->str ? V(v)->str : ""; }
void* fk_buffer_data(void* v) { return V(v)->data; }
size_t fk_buffer_len(void* v) { return V(v)->len; }
const char* fk_exception(void) { return g_exception; }
void fk_clear_exception(void) { g_exception[0] = 0; }

napi_value napi_register_module_v1(napi_env env, napi_value exports);
void* fk_init_module(void) { fk_value* e = mk(FK_OBJECT); return napi_register_module_v1((napi_env)fk_env(), N(e)); }

static void* call(fk_value* fn_self, napi_callback cb, size_t argc, void** argv) {
    struct napi_callback_info__ info = {argc, (fk_value**)argv, fn_self};
    g_exception[0] = 0;
    napi_value r = cb((napi_env)fk_env(), &info);
    return g_exception[0] ? NULL : r;
}
void* fk_new(void* cls, size_t argc, void** argv) {
    fk_value* inst = mk(FK_OBJECT);
    inst->cls = V(cls);
    return call(inst, V(cls)->cb, argc, argv) ? inst : NULL;
}
void* fk_call_function(void* fn, size_t argc, void** argv) {
    if (!fn || V(fn)->kind != FK_FUNCTION) { snprintf(g_exception, sizeof g_exception, "not a function"); return NULL; }
    return call(NULL, V(fn)->cb, argc, argv);
}
void* fk_call_method(void* inst, const char* name, size_t argc, void** argv) {
    fk_value* m = V(inst)->cls ? prop_get(V(inst)->cls, name) : NULL;
    if (!m) { snprintf(g_exception, sizeof g_exception, "no method %s", name); return NULL; }
    return call(V(inst), m->cb, argc, argv);
}

/* ---------------- Node-API doubles ---------------- */
napi_status napi_define_class(napi_env env, const char* name, size_t length, napi_callback ctor, void* data, size_t n,
                              const napi_property_descriptor* props, napi_value* result) {
    (void)env; (void)length; (void)data;
    fk_value* c = mk(FK_CLASS);
    c->str = strdup(name); c->cb = ctor;
    for (size_t i = 0; i < n; ++i) { fk_value* f = mk(FK_FUNCTION); f->cb = props[i].method; prop_set(c, props[i].utf8name, f); }
    *result = N(c);
    return napi_ok;
}
napi_status napi_set_named_property(napi_env env, napi_value o, const char* name, napi_value v) { (void)env; prop_set(V(o), name, V(v)); return napi_ok; }
napi_status napi_get_named_property(napi_env env, napi_value o, const char* name, napi_value* r) {
    (void)env; fk_value* v = prop_get(V(o), name); *r = N(v ? v : mk(FK_UNDEF)); return napi_ok;
}
napi_status napi_has_named_property(napi_env env, napi_value o, const char* name, bool* r) {
    (void)env; if (V(o)->kind != FK_OBJECT) return napi_object_expected; *r = prop_get(V(o), name) != NULL; return napi_ok;
}
napi_status napi_get_cb_info(napi_env env, napi_callback_info info, size_t* argc, napi_value* argv, napi_value* self, void** data) {
    (void)env;
    if (argc) {
        size_t want = *argc;
        for (size_t i = 0; i < want && argv; ++i) argv[i] = N(i < info->argc ? info->argv[i] : mk(FK_UNDEF));
        *argc = info->argc;
    }
    if (self) *self = N(info->self);
    if (data) *data = NULL;
    return napi_ok;
}
napi_status napi_wrap(napi_env env, napi_value o, void* native, napi_finalize fin, void* hint, napi_ref* r) {
    (void)env; (void)fin; (void)hint; (void)r; V(o)->wrapped = native; return napi_ok;
}
napi_status napi_unwrap(napi_env env, napi_value o, void** r) { (void)env; *r = V(o)->wrapped; return *r ? napi_ok : napi_invalid_arg; }
napi_status napi_typeof(napi_env env, napi_value v, napi_valuetype* r) {
    (void)env;
    switch (V(v)->kind) {
        case FK_NUMBER: *r = napi_number; break;
        case FK_STRING: *r = napi_string; break;
        case FK_UNDEF: *r = napi_undefined; break;
        case FK_FUNCTION: case FK_CLASS: *r = napi_function; break;
        default: *r = napi_object;
    }
    return napi_ok;
}
napi_status napi_get_value_double(napi_env env, napi_value v, double* r) { (void)env; if (V(v)->kind != FK_NUMBER) return napi_number_expected; *r = V(v)->num; return napi_ok; }
napi_status napi_get_value_int64(napi_env env, napi_value v, int64_t* r) { (void)env; if (V(v)->kind != FK_NUMBER) return napi_number_expected; *r = (int64_t)V(v)->num; return napi_ok; }
napi_status napi_get_value_uint32(napi_env env, napi_value v, uint32_t* r) { (void)env; if (V(v)->kind != FK_NUMBER) return napi_number_expected; *r = (uint32_t)V(v)->num; return napi_ok; }
napi_status napi_get_value_string_utf8(napi_env env, napi_value v, char* buf, size_t n, size_t* r) {
    (void)env; if (V(v)->kind != FK_STRING) return napi_string_expected;
    snprintf(buf, n, "%s", V(v)->str); if (r) *r = strlen(buf); return napi_ok;
}
napi_status napi_is_array(napi_env env, napi_value v, bool* r) { (void)env; *r = V(v)->kind == FK_ARRAY; return napi_ok; }
napi_status napi_get_array_length(napi_env env, napi_value v, uint32_t* r) { (void)env; if (V(v)->kind != FK_ARRAY) return napi_array_expected; *r = V(v)->nelems; return napi_ok; }
napi_status napi_get_element(napi_env env, napi_value v, uint32_t i, napi_value* r) { (void)env; *r = N(i < V(v)->nelems ? V(v)->elems[i] : mk(FK_UNDEF)); return napi_ok; }
napi_status napi_get_typedarray_info(napi_env env, napi_value v, napi_typedarray_type* t, size_t* len, void** data, napi_value* ab, size_t* off) {
    (void)env; if (V(v)->kind != FK_TYPED) return napi_invalid_arg;
    if (t) *t = V(v)->ttype; if (len) *len = V(v)->len; if (data) *data = V(v)->data; if (ab) *ab = NULL; if (off) *off = 0;
    return napi_ok;
}
napi_status napi_create_arraybuffer(napi_env env, size_t n, void** data, napi_value* r) {
    (void)env; fk_value* b = mk(FK_BUFFER); b->data = calloc(n ? n : 1, 1); b->len = n; if (data) *data = b->data; *r = N(b); return napi_ok;
}
napi_status napi_create_object(napi_env env, napi_value* r) { (void)env; *r = N(mk(FK_OBJECT)); return napi_ok; }
napi_status napi_create_double(napi_env env, double d, napi_value* r) { (void)env; *r = N(fk_number(d)); return napi_ok; }
napi_status napi_create_int64(napi_env env, int64_t d, napi_value* r) { (void)env; *r = N(fk_number((double)d)); return napi_ok; }
napi_status napi_create_string_utf8(napi_env env, const char* s, size_t n, napi_value* r) { (void)env; (void)n; *r = N(fk_string(s)); return napi_ok; }
napi_status napi_get_undefined(napi_env env, napi_value* r) { (void)env; *r = N(mk(FK_UNDEF)); return napi_ok; }
napi_status napi_create_function(napi_env env, const char* name, size_t len, napi_callback cb, void* data, napi_value* r) {
    (void)env; (void)name; (void)len; (void)data; fk_value* f = mk(FK_FUNCTION); f->cb = cb; *r = N(f); return napi_ok;
}
napi_status napi_create_external_arraybuffer(napi_env env, void* data, size_t n, napi_finalize fin, void* hint, napi_value* r) {
    (void)env; (void)fin; (void)hint; fk_value* b = mk(FK_BUFFER); b->data = data; b->len = n; *r = N(b); return napi_ok;
}
napi_status napi_throw_error(napi_env env, const char* code, const char* msg) {
    (void)env; snprintf(g_exception, sizeof g_exception, "%s: %s", code ? code : "", msg ? msg : ""); return napi_ok;
}
