/*
 * dusp_b200_addon.c -- Node.js N-API addon over the C ABI (include/dusp_b200.h).
 *
 * NOT BUILT IN THIS IMAGE: node, npm and node_api.h are absent from the build
 * container and from the GPU box (probed: `node --version` -> command not
 * found).  The file is the binding a maintainer of the reference adds; it only
 * uses the ABI-stable N-API C interface and the six entry points below.  The
 * Python host mirror (dusp_b200/capi.py) exercises exactly the same calls.
 *
 *   renderBatch(blob: string, nVoices: number, paramNames: string[],
 *               paramValues: Float64Array /* [nVoices][paramNames.length] *\/,
 *               nSamples: number, outMode: 0|1) -> Promise<Float32Array>
 *
 * The render runs in napi_create_async_work so the JS signature stays async
 * like the reference's renderChannelData (src/renderChannelData.js:5); errors
 * reject with dusp_last_error() as a *string*, the reference's convention
 * (renderChannelData.js:13,17).
 */
#include <node_api.h>
#include <stdlib.h>
#include <string.h>

#include "../../include/dusp_b200.h"

typedef struct {
  napi_async_work work;
  napi_deferred deferred;
  char *blob;
  size_t blob_len;
  int n_voices, n_params, out_mode;
  char **names;
  double *values;
  long long n_samples;
  float *out;
  size_t out_floats;
  int rc;
  char err[512];
} Job;

static void execute(napi_env env, void *data) {
  Job *j = (Job *)data;
  dusp_program *p = NULL;
  dusp_session *s = NULL;
  j->rc = dusp_program_create(j->blob, j->blob_len, &p);
  if (j->rc == DUSP_OK) {
    const dusp_program *pp = p;
    j->rc = dusp_session_create(&pp, &j->n_voices, 1, 0, &s);
  }
  if (j->rc == DUSP_OK && j->n_params > 0) {
    int *ids = (int *)malloc(sizeof(int) * j->n_params);
    for (int k = 0; k < j->n_params; k++) ids[k] = dusp_program_param_index(p, j->names[k]);
    j->rc = dusp_session_set_params(s, 0, ids, j->n_params, j->values);
    free(ids);
  }
  if (j->rc == DUSP_OK) {
    int nch = j->out_mode == DUSP_OUT_MIX ? 2 : dusp_session_num_channels(s);
    size_t rows = j->out_mode == DUSP_OUT_MIX ? 2 : (size_t)j->n_voices * nch;
    j->out_floats = rows * (size_t)j->n_samples;
    j->out = (float *)malloc(j->out_floats * sizeof(float) + 16);
    j->rc = dusp_session_render(s, j->n_samples, j->out_mode, j->out, NULL, NULL);
  }
  if (j->rc != DUSP_OK) strncpy(j->err, dusp_last_error(), sizeof j->err - 1);
  if (s) dusp_session_destroy(s);
  if (p) dusp_program_destroy(p);
}

static void free_out(napi_env env, void *data, void *hint) { free(data); }

static void complete(napi_env env, napi_status status, void *data) {
  Job *j = (Job *)data;
  if (j->rc == DUSP_OK) {
    napi_value ab, ta;
    napi_create_external_arraybuffer(env, j->out, j->out_floats * sizeof(float), free_out, NULL, &ab);
    napi_create_typedarray(env, napi_float32_array, j->out_floats, ab, 0, &ta);
    napi_resolve_deferred(env, j->deferred, ta);
  } else {
    napi_value msg;
    napi_create_string_utf8(env, j->err, NAPI_AUTO_LENGTH, &msg);  /* reject with a string, like `throw "..."` */
    napi_reject_deferred(env, j->deferred, msg);
    free(j->out);
  }
  napi_delete_async_work(env, j->work);
  for (int k = 0; k < j->n_params; k++) free(j->names[k]);
  free(j->names); free(j->values); free(j->blob); free(j);
}

static napi_value render_batch(napi_env env, napi_callback_info info) {
  size_t argc = 6;
  napi_value argv[6], promise, name;
  napi_get_cb_info(env, info, &argc, argv, NULL, NULL);
  Job *j = (Job *)calloc(1, sizeof(Job));
  napi_get_value_string_utf8(env, argv[0], NULL, 0, &j->blob_len);
  j->blob = (char *)malloc(j->blob_len + 1);
  napi_get_value_string_utf8(env, argv[0], j->blob, j->blob_len + 1, &j->blob_len);
  napi_get_value_int32(env, argv[1], &j->n_voices);
  uint32_t np = 0;
  napi_get_array_length(env, argv[2], &np);
  j->n_params = (int)np;
  j->names = (char **)calloc(np ? np : 1, sizeof(char *));
  for (uint32_t k = 0; k < np; k++) {
    napi_value e; size_t len;
    napi_get_element(env, argv[2], k, &e);
    napi_get_value_string_utf8(env, e, NULL, 0, &len);
    j->names[k] = (char *)malloc(len + 1);
    napi_get_value_string_utf8(env, e, j->names[k], len + 1, &len);
  }
  if (np) {
    void *data; size_t len; napi_typedarray_type tt; napi_value ab; size_t off;
    napi_get_typedarray_info(env, argv[3], &tt, &len, &data, &ab, &off);
    j->values = (double *)malloc(len * sizeof(double));
    memcpy(j->values, data, len * sizeof(double));
  }
  int64_t ns; napi_get_value_int64(env, argv[4], &ns); j->n_samples = ns;
  napi_get_value_int32(env, argv[5], &j->out_mode);
  napi_create_promise(env, &j->deferred, &promise);
  napi_create_string_utf8(env, "dusp_b200.renderBatch", NAPI_AUTO_LENGTH, &name);
  napi_create_async_work(env, NULL, name, execute, complete, j, &j->work);
  napi_queue_async_work(env, j->work);
  return promise;
}

static napi_value init(napi_env env, napi_value exports) {
  napi_value fn;
  napi_create_function(env, "renderBatch", NAPI_AUTO_LENGTH, render_batch, NULL, &fn);
  napi_set_named_property(env, exports, "renderBatch", fn);
  return exports;
}
NAPI_MODULE(NODE_GYP_MODULE_NAME, init)
