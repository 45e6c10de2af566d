/* kmeans_napi.c — Node-API addon over libkmeans.so.
 *
 * STATUS: built, unexecuted.  There is no Node.js (and no <node_api.h>) on this image, so this
 * file is compiled against napi/node_api_min.h in tests/test_napi_shim.py to prove it is
 * well-formed and that every km_* call matches include/kmeans.h; it has never run under Node.
 *
 * What it binds: the four entry points the reference performs by hand —
 *   init    /root/reference/app.mjs:126-129 (addCentroid), :143-145 (addCard)
 *   assign  /root/reference/app.mjs:146-156 (updateCardAssign)
 *   update  /root/reference/app.mjs:450-461, 476-480, 571-573 (summarise + rename a centroid)
 *   run     /root/reference/app.mjs:286-288, 498-508 (the #iter counter)
 * Typed arrays are borrowed for the duration of one call; results are copied into arrays the
 * caller allocates, so nothing here needs a finalizer except the context itself.
 */
#ifdef KM_NAPI_SELF_DECLARED
#include "node_api_min.h"
#else
#include <node_api.h>
#endif
#include "kmeans.h"

#define KM_NAPI_MAX_ARGS 6

static napi_value km_undefined(napi_env env) {
  napi_value u;
  napi_get_undefined(env, &u);
  return u;
}

/* non-zero status -> thrown JS Error carrying km_last_error() */
static napi_value km_throw(napi_env env, km_ctx *ctx) {
  napi_throw_error(env, "ERR_KMEANS", km_last_error(ctx));
  return NULL;
}

static void km_finalize_ctx(napi_env env, void *data, void *hint) {
  (void)env;
  (void)hint;
  if (data) km_destroy((km_ctx *)data);
}

static km_ctx *km_unwrap(napi_env env, napi_value v) {
  void *p = NULL;
  if (napi_get_value_external(env, v, &p) != napi_ok || !p) {
    napi_throw_error(env, "ERR_KMEANS", "first argument must be a state returned by init()");
    return NULL;
  }
  return (km_ctx *)p;
}

static napi_value km_stats_object(napi_env env, const km_stats *s) {
  napi_value o, v;
  napi_create_object(env, &o);
  napi_create_int32(env, s->iterations, &v);       napi_set_named_property(env, o, "iterations", v);
  napi_get_boolean(env, s->converged != 0, &v);    napi_set_named_property(env, o, "converged", v);
  napi_create_int64(env, s->changed, &v);          napi_set_named_property(env, o, "changed", v);
  napi_create_double(env, s->inertia, &v);         napi_set_named_property(env, o, "inertia", v);
  napi_create_int64(env, s->rescored, &v);         napi_set_named_property(env, o, "rescored", v);
  napi_create_double(env, (double)s->device_ms, &v); napi_set_named_property(env, o, "deviceMs", v);
  return o;
}

/* init(points: Float32Array|Float64Array, d, k, centroids: Float64Array | null, seed) -> state */
static napi_value Init(napi_env env, napi_callback_info info) {
  size_t argc = 5;
  napi_value argv[KM_NAPI_MAX_ARGS];
  napi_get_cb_info(env, info, &argc, argv, NULL, NULL);
  if (argc < 5) { napi_throw_error(env, "ERR_KMEANS", "init(points, d, k, centroids|null, seed)"); return NULL; }
  napi_typedarray_type ty;
  size_t len = 0;
  void *pts = NULL;
  if (napi_get_typedarray_info(env, argv[0], &ty, &len, &pts, NULL, NULL) != napi_ok ||
      (ty != napi_float32_array && ty != napi_float64_array)) {
    napi_throw_error(env, "ERR_KMEANS", "points must be a Float32Array or Float64Array");
    return NULL;
  }
  int32_t d = 0, k = 0;
  int64_t seed = 0;
  napi_get_value_int32(env, argv[1], &d);
  napi_get_value_int32(env, argv[2], &k);
  napi_get_value_int64(env, argv[4], &seed);
  if (d < 1 || len % (size_t)d != 0) { napi_throw_error(env, "ERR_KMEANS", "points.length must be a multiple of d"); return NULL; }

  km_config cfg = {sizeof(km_config), 0, KM_PATH_AUTO, KM_UPDATE_INCREMENTAL, 1, 0, NULL, 0, 0};
  km_ctx *ctx = NULL;
  if (km_create(&cfg, &ctx) != KM_OK) return km_throw(env, NULL);
  km_status st = km_upload_points(ctx, pts, (int64_t)(len / (size_t)d), d,
                                  ty == napi_float64_array ? KM_DTYPE_F64 : KM_DTYPE_F32);
  if (st == KM_OK) {
    napi_valuetype vt;
    napi_typeof(env, argv[3], &vt);
    if (vt == napi_null || vt == napi_undefined) {
      st = km_init_forgy(ctx, k, (uint64_t)seed);   /* the seeded form of app.mjs:163 */
    } else {
      napi_typedarray_type cty;
      size_t clen = 0;
      void *c0 = NULL;
      if (napi_get_typedarray_info(env, argv[3], &cty, &clen, &c0, NULL, NULL) != napi_ok ||
          cty != napi_float64_array || clen != (size_t)k * (size_t)d) {
        km_destroy(ctx);
        napi_throw_error(env, "ERR_KMEANS", "centroids must be a Float64Array of k*d values");
        return NULL;
      }
      st = km_set_centroids(ctx, (const double *)c0, k);
    }
  }
  if (st != KM_OK) {
    napi_value r = km_throw(env, ctx);
    km_destroy(ctx);
    return r;
  }
  napi_value ext;
  napi_create_external(env, ctx, km_finalize_ctx, NULL, &ext);
  return ext;
}

typedef km_status (*km_stats_fn)(km_ctx *, km_stats *);
static napi_value km_call_stats(napi_env env, napi_callback_info info, km_stats_fn fn) {
  size_t argc = 1;
  napi_value argv[KM_NAPI_MAX_ARGS];
  napi_get_cb_info(env, info, &argc, argv, NULL, NULL);
  km_ctx *ctx = argc ? km_unwrap(env, argv[0]) : NULL;
  if (!ctx) return NULL;
  km_stats s;
  if (fn(ctx, &s) != KM_OK) return km_throw(env, ctx);
  return km_stats_object(env, &s);
}
static napi_value Assign(napi_env env, napi_callback_info info) { return km_call_stats(env, info, km_assign); }
static napi_value Update(napi_env env, napi_callback_info info) { return km_call_stats(env, info, km_update); }
static napi_value Step(napi_env env, napi_callback_info info) { return km_call_stats(env, info, km_step); }

/* run(state, maxIter) -> stats */
static napi_value Run(napi_env env, napi_callback_info info) {
  size_t argc = 2;
  napi_value argv[KM_NAPI_MAX_ARGS];
  napi_get_cb_info(env, info, &argc, argv, NULL, NULL);
  km_ctx *ctx = argc ? km_unwrap(env, argv[0]) : NULL;
  if (!ctx) return NULL;
  int32_t max_iter = 300;
  if (argc > 1) napi_get_value_int32(env, argv[1], &max_iter);
  km_stats s;
  if (km_run(ctx, max_iter, &s) != KM_OK) return km_throw(env, ctx);
  return km_stats_object(env, &s);
}

/* labels(state, out: Int32Array | Uint8Array), centroids(state, out: Float64Array), counts(state, out: BigInt64Array) */
static napi_value km_copy_out(napi_env env, napi_callback_info info, int what) {
  size_t argc = 2;
  napi_value argv[KM_NAPI_MAX_ARGS];
  napi_get_cb_info(env, info, &argc, argv, NULL, NULL);
  km_ctx *ctx = argc ? km_unwrap(env, argv[0]) : NULL;
  if (!ctx) return NULL;
  napi_typedarray_type ty;
  size_t len = 0;
  void *out = NULL;
  if (argc < 2 || napi_get_typedarray_info(env, argv[1], &ty, &len, &out, NULL, NULL) != napi_ok) {
    napi_throw_error(env, "ERR_KMEANS", "second argument must be the output typed array");
    return NULL;
  }
  km_status st = KM_ERR_INVALID_ARG;
  int64_t n_local = 0;
  int32_t d = 0, k = 0;
  km_get_shape(ctx, &n_local, NULL, &d, &k);
  /* the C calls write n, k*d and k elements: a shorter typed array would be written past its end */
  const size_t need = what == 0 ? (size_t)n_local : what == 1 ? (size_t)k * (size_t)d : (size_t)k;
  if (len < need) {
    napi_throw_error(env, "ERR_KMEANS", "output typed array is shorter than the result (labels: n, centroids: k*d, counts: k)");
    return NULL;
  }
  if (what == 0 && ty == napi_int32_array) st = km_get_labels(ctx, (int32_t *)out, n_local);
  else if (what == 0 && ty == napi_uint8_array) st = km_get_labels_u8(ctx, (uint8_t *)out, n_local);   /* k <= 255: a byte per label */
  else if (what == 1 && ty == napi_float64_array) st = km_get_centroids(ctx, (double *)out);
  else if (what == 2 && ty == napi_bigint64_array) st = km_get_counts(ctx, (int64_t *)out);
  else { napi_throw_error(env, "ERR_KMEANS", "output array has the wrong element type"); return NULL; }
  if (st != KM_OK) return km_throw(env, ctx);
  return km_undefined(env);
}
static napi_value Labels(napi_env env, napi_callback_info info) { return km_copy_out(env, info, 0); }
static napi_value Centroids(napi_env env, napi_callback_info info) { return km_copy_out(env, info, 1); }
static napi_value Counts(napi_env env, napi_callback_info info) { return km_copy_out(env, info, 2); }

/* module entry point: the symbol Node looks up in a .node file (NAPI_MODULE_INIT expands to it) */
#if defined(__GNUC__)
__attribute__((visibility("default")))
#endif
napi_value napi_register_module_v1(napi_env env, napi_value exports) {
  const napi_property_descriptor props[] = {
      {"init", NULL, Init, NULL, NULL, NULL, napi_default, NULL},
      {"assign", NULL, Assign, NULL, NULL, NULL, napi_default, NULL},
      {"update", NULL, Update, NULL, NULL, NULL, napi_default, NULL},
      {"step", NULL, Step, NULL, NULL, NULL, napi_default, NULL},
      {"run", NULL, Run, NULL, NULL, NULL, napi_default, NULL},
      {"labels", NULL, Labels, NULL, NULL, NULL, napi_default, NULL},
      {"centroids", NULL, Centroids, NULL, NULL, NULL, napi_default, NULL},
      {"counts", NULL, Counts, NULL, NULL, NULL, napi_default, NULL},
  };
  napi_define_properties(env, exports, sizeof(props) / sizeof(props[0]), props);
  return exports;
}
