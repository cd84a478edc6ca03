/*
 * dusp_b200_addon.c -- Node.js N-API addon over the C ABI (include/dusp_b200.h).
 *
 * NOT BUILT IN THIS IMAGE: node, npm and node_api.h are absent from the build container and from the GPU box
 * (probed: `node --version` -> command not found).  The file is the binding a maintainer of the reference adds; it
 * uses only the ABI-stable N-API C interface and five entry points of the library: dusp_run_create,
 * dusp_run_set_params, dusp_run_num_channels / dusp_run_channels_rendered, dusp_run_render, dusp_run_destroy.  The
 * Python host mirror (dusp_b200/capi.py: Run) makes exactly the same calls.
 *
 *   renderRun(blobs: string[], voices: number[], paramNames: string[][], paramValues: Float64Array[],
 *             nSamples: number, outMode: 0|1|2, device: number)
 *       -> Promise<{ data: Float32Array, channels: number, voices: number }>
 *
 * blobs[i] is the "DUSPGRAPH 1" text of program group i (with its EV / TG / EPOCH / KEEP records, index.js),
 * instantiated voices[i] times; paramValues[i] is [voices[i]][paramNames[i].length] row-major.  `data` is
 * [voice][channel][nSamples] (outMode 0 / 2) or [2][nSamples] (outMode 1, stereo mixdown).
 *
 * The render runs in napi_create_async_work so the JS signature stays async like the reference's renderChannelData
 * (src/renderChannelData.js:5); errors reject with dusp_last_error() as a *string*, the reference's convention
 * (renderChannelData.js:13,17).
 */
#include <node_api.h>
#include <stdlib.h>
#include <string.h>

#include "../../include/dusp_b200.h"

typedef struct {
  napi_async_work work;
  napi_deferred deferred;
  int n_programs, out_mode, device;
  char **blobs;
  size_t *blob_lens;
  int *voices;
  int *n_names;
  char ***names;    /* [program][name] */
  double **values;  /* [program] -> [voices][n_names] */
  long long n_samples;
  float *out;
  size_t out_floats;
  int channels, total_voices;
  int rc;
  char err[1024];
} Job;

static void job_free(Job *j) {
  if (!j) return;
  for (int i = 0; i < j->n_programs; i++) {
    if (j->blobs) free(j->blobs[i]);
    if (j->names && j->names[i]) { for (int k = 0; k < j->n_names[i]; k++) free(j->names[i][k]); free(j->names[i]); }
    if (j->values) free(j->values[i]);
  }
  free(j->blobs); free(j->blob_lens); free(j->voices); free(j->n_names); free(j->names); free(j->values);
  free(j);
}

static void execute(napi_env env, void *data) {
  Job *j = (Job *)data;
  dusp_run *r = NULL;
  (void)env;
  j->rc = dusp_run_create((const char *const *)j->blobs, j->blob_lens, j->voices, j->n_programs, j->device, &r);
  for (int i = 0; j->rc == DUSP_OK && i < j->n_programs; i++)
    if (j->n_names[i] > 0)
      j->rc = dusp_run_set_params(r, i, (const char *const *)j->names[i], j->n_names[i], j->values[i]);
  if (j->rc == DUSP_OK) {
    const int nch = dusp_run_num_channels(r);
    j->total_voices = dusp_run_num_voices(r);
    const size_t rows = j->out_mode == DUSP_OUT_MIX ? 2 : (size_t)j->total_voices * (size_t)nch;
    j->out_floats = rows * (size_t)j->n_samples;
    j->out = (float *)malloc(j->out_floats * sizeof(float) + 16);
    if (!j->out) { j->rc = DUSP_ENOMEM; strncpy(j->err, "out of host memory for the rendered channel data", sizeof j->err - 1); }
    else j->rc = dusp_run_render(r, j->n_samples, j->out_mode, j->out, NULL, NULL);
    /* channels the reference's result would hold (renderChannelData.js:35-45 grows them as they appear); rows are
     * laid out with the widest epoch's count, so the caller strides by `rowChannels` and keeps `channels` of them */
    j->channels = j->out_mode == DUSP_OUT_MIX ? 2 : (j->n_samples ? dusp_run_channels_rendered(r) : 0);
    if (j->rc == DUSP_OK && j->out_mode != DUSP_OUT_MIX && j->channels < nch && j->n_samples) {
      /* compact [voice][nch][n] -> [voice][channels][n] in place (rows only move towards the front) */
      for (int v = 0; v < j->total_voices; v++)
        memmove(j->out + (size_t)v * j->channels * j->n_samples, j->out + (size_t)v * nch * j->n_samples,
                (size_t)j->channels * j->n_samples * sizeof(float));
      j->out_floats = (size_t)j->total_voices * j->channels * (size_t)j->n_samples;
    }
  }
  if (j->rc != DUSP_OK && !j->err[0]) strncpy(j->err, dusp_last_error(), sizeof j->err - 1);
  if (r) dusp_run_destroy(r);
}

static void free_out(napi_env env, void *data, void *hint) { (void)env; (void)hint; free(data); }

static void complete(napi_env env, napi_status status, void *data) {
  Job *j = (Job *)data;
  if (status == napi_ok && j->rc == DUSP_OK) {
    napi_value ab, ta, obj, v;
    napi_create_external_arraybuffer(env, j->out, j->out_floats * sizeof(float), free_out, NULL, &ab);
    napi_create_typedarray(env, napi_float32_array, j->out_floats, ab, 0, &ta);
    napi_create_object(env, &obj);
    napi_set_named_property(env, obj, "data", ta);
    napi_create_int32(env, j->channels, &v); napi_set_named_property(env, obj, "channels", v);
    napi_create_int32(env, j->total_voices, &v); napi_set_named_property(env, obj, "voices", v);
    napi_resolve_deferred(env, j->deferred, obj);
  } else {
    napi_value msg;
    napi_create_string_utf8(env, j->err[0] ? j->err : "dusp_b200: render cancelled", NAPI_AUTO_LENGTH, &msg);
    napi_reject_deferred(env, j->deferred, msg);  /* reject with a string, like `throw "..."` */
    free(j->out);
  }
  napi_delete_async_work(env, j->work);
  job_free(j);
}

#define NAPI_TRY(call, what) do { if ((call) != napi_ok) { napi_throw_type_error(env, NULL, "renderRun: " what); job_free(j); return NULL; } } while (0)

static char *get_string(napi_env env, napi_value v, size_t *len_out) {
  size_t len = 0;
  if (napi_get_value_string_utf8(env, v, NULL, 0, &len) != napi_ok) return NULL;
  char *s = (char *)malloc(len + 1);
  if (!s) return NULL;
  if (napi_get_value_string_utf8(env, v, s, len + 1, &len) != napi_ok) { free(s); return NULL; }
  if (len_out) *len_out = len;
  return s;
}

static napi_value render_run(napi_env env, napi_callback_info info) {
  size_t argc = 7;
  napi_value argv[7], promise, name;
  Job *j = (Job *)calloc(1, sizeof(Job));
  if (!j) return NULL;
  NAPI_TRY(napi_get_cb_info(env, info, &argc, argv, NULL, NULL), "bad call");
  if (argc < 6) { napi_throw_type_error(env, NULL, "renderRun(blobs, voices, paramNames, paramValues, nSamples, outMode[, device])"); job_free(j); return NULL; }
  uint32_t np = 0, nv = 0, nn = 0, nvals = 0;
  NAPI_TRY(napi_get_array_length(env, argv[0], &np), "blobs must be an array of strings");
  NAPI_TRY(napi_get_array_length(env, argv[1], &nv), "voices must be an array of numbers");
  NAPI_TRY(napi_get_array_length(env, argv[2], &nn), "paramNames must be an array (one array of names per program, or empty)");
  NAPI_TRY(napi_get_array_length(env, argv[3], &nvals), "paramValues must be an array (one Float64Array per program, or empty)");
  if (np < 1 || nv != np || (nn != 0 && nn != np) || nvals != nn) { napi_throw_type_error(env, NULL, "renderRun: blobs, voices, paramNames and paramValues must describe the same programs"); job_free(j); return NULL; }
  j->n_programs = (int)np;
  j->blobs = (char **)calloc(np, sizeof(char *)); j->blob_lens = (size_t *)calloc(np, sizeof(size_t));
  j->voices = (int *)calloc(np, sizeof(int)); j->n_names = (int *)calloc(np, sizeof(int));
  j->names = (char ***)calloc(np, sizeof(char **)); j->values = (double **)calloc(np, sizeof(double *));
  if (!j->blobs || !j->blob_lens || !j->voices || !j->n_names || !j->names || !j->values) { napi_throw_error(env, NULL, "renderRun: out of memory"); job_free(j); return NULL; }
  for (uint32_t i = 0; i < np; i++) {
    napi_value e;
    NAPI_TRY(napi_get_element(env, argv[0], i, &e), "blobs");
    j->blobs[i] = get_string(env, e, &j->blob_lens[i]);
    if (!j->blobs[i]) { napi_throw_type_error(env, NULL, "renderRun: blobs must be strings"); job_free(j); return NULL; }
    NAPI_TRY(napi_get_element(env, argv[1], i, &e), "voices");
    NAPI_TRY(napi_get_value_int32(env, e, &j->voices[i]), "voices must be numbers");
    if (j->voices[i] < 1) { napi_throw_range_error(env, NULL, "renderRun: every program needs at least one voice"); job_free(j); return NULL; }
    if (nn == 0) continue;
    napi_value names_i, vals_i;
    uint32_t k_names = 0;
    NAPI_TRY(napi_get_element(env, argv[2], i, &names_i), "paramNames");
    NAPI_TRY(napi_get_array_length(env, names_i, &k_names), "paramNames[i] must be an array of strings");
    j->n_names[i] = (int)k_names;
    j->names[i] = (char **)calloc(k_names ? k_names : 1, sizeof(char *));
    for (uint32_t k = 0; k < k_names; k++) {
      NAPI_TRY(napi_get_element(env, names_i, k, &e), "paramNames");
      j->names[i][k] = get_string(env, e, NULL);
      if (!j->names[i][k]) { napi_throw_type_error(env, NULL, "renderRun: parameter names must be strings"); job_free(j); return NULL; }
    }
    if (k_names) {
      void *data; size_t len; napi_typedarray_type tt; napi_value ab; size_t off;
      NAPI_TRY(napi_get_element(env, argv[3], i, &vals_i), "paramValues");
      NAPI_TRY(napi_get_typedarray_info(env, vals_i, &tt, &len, &data, &ab, &off), "paramValues[i] must be a Float64Array");
      if (tt != napi_float64_array || len != (size_t)j->voices[i] * k_names) {
        napi_throw_type_error(env, NULL, "renderRun: paramValues[i] must be a Float64Array of voices[i] * paramNames[i].length values");
        job_free(j); return NULL;
      }
      j->values[i] = (double *)malloc(len * sizeof(double));
      if (!j->values[i]) { napi_throw_error(env, NULL, "renderRun: out of memory"); job_free(j); return NULL; }
      memcpy(j->values[i], data, len * sizeof(double));
    }
  }
  int64_t ns = 0;
  NAPI_TRY(napi_get_value_int64(env, argv[4], &ns), "nSamples must be a number");
  NAPI_TRY(napi_get_value_int32(env, argv[5], &j->out_mode), "outMode must be 0, 1 or 2");
  if (argc > 6) NAPI_TRY(napi_get_value_int32(env, argv[6], &j->device), "device must be a number");
  if (ns < 0 || j->out_mode < 0 || j->out_mode > 2) { napi_throw_range_error(env, NULL, "renderRun: nSamples >= 0 and outMode in 0..2"); job_free(j); return NULL; }
  j->n_samples = ns;
  NAPI_TRY(napi_create_promise(env, &j->deferred, &promise), "promise");
  NAPI_TRY(napi_create_string_utf8(env, "dusp_b200.renderRun", NAPI_AUTO_LENGTH, &name), "name");
  NAPI_TRY(napi_create_async_work(env, NULL, name, execute, complete, j, &j->work), "async work");
  NAPI_TRY(napi_queue_async_work(env, j->work), "queue");
  return promise;
}

static napi_value init(napi_env env, napi_value exports) {
  napi_value fn;
  napi_create_function(env, "renderRun", NAPI_AUTO_LENGTH, render_run, NULL, &fn);
  napi_set_named_property(env, exports, "renderRun", fn);
  return exports;
}
NAPI_MODULE(NODE_GYP_MODULE_NAME, init)
