/*
 * rt_addon.c -- Node-API addon over the C-ABI of include/rt_engine.h: the in-process route from the reference's
 * TypeScript to the B200 engine.
 *
 * The reference reaches its local model through fetch() (src/adapters/local-llm.ts:112,159); with this addon the
 * adapter class native/local-llm-b200.ts calls the engine directly.  rt_generate() blocks for the whole generation, so
 * it runs on the libuv pool (napi_async_work) and settles a Promise -- execute() stays `Promise<string>`
 * (src/adapters/base.ts:16).
 *
 * Exports:  create(model: string, device: number, maxCtx: number, maxNewTokens?: number) -> engine (external)
 *           generate(engine, prompt: string, timeoutMs: number) -> Promise<string>
 *           contextLength(engine) -> number           (rt_model_info_get; feeds getMaxSourceChars)
 *           destroy(engine) -> undefined
 *
 * Build where Node >= 20 exists (package.json:53-55 of the reference):
 *   gcc -shared -fPIC native/rt_addon.c -I"$(node -p 'process.execPath')/../../include/node" -Iinclude \
 *       -Ltheroundtaible_b200/_lib -lrt_engine -o native/rt_addon.node
 * NOT BUILT on this image (no node, no node_api.h); tests/test_native_sources.py only checks that it compiles against a
 * declarations-only stand-in of the N-API functions it uses.
 */
#include <node_api.h>
#include <stdint.h>
#include <stdlib.h>
#include <string.h>

#include "rt_engine.h"

typedef struct {
    rt_engine* engine;
    int32_t max_new_tokens;
} addon_engine;

typedef struct {
    napi_async_work work;
    napi_deferred deferred;
    addon_engine* ae;
    char* prompt;
    size_t len;
    uint64_t deadline_ms;
    rt_result res;
    int rc;
    char err[1024];
} job_t;

static napi_value throw_text(napi_env env, const char* text) {
    napi_throw_error(env, NULL, text);
    return NULL;
}

/* worker thread: no N-API calls here */
static void job_run(napi_env env, void* data) {
    job_t* j = (job_t*)data;
    rt_sampling s;
    (void)env;
    memset(&s, 0, sizeof s);                        /* greedy, stop at <eos>; the reference request carries no sampling field */
    s.max_new_tokens = j->ae->max_new_tokens;       /* 0 -> the engine's default of 512 */
    j->rc = rt_generate(j->ae->engine, j->prompt, j->len, &s, j->deadline_ms, &j->res);
    if (j->rc != 0) {
        strncpy(j->err, rt_last_error(j->ae->engine), sizeof j->err - 1);
        j->err[sizeof j->err - 1] = 0;
    }
}

/* main thread: settle the Promise.  The rejection text is rt_last_error(): "... timed out ..." for a passed deadline,
 * "context length exceeded ..." for a prompt that does not fit -- what classifyError (src/utils/errors.ts:86-126) and
 * isContextWindowError (local-llm.ts:240-248) look for. */
static void job_done(napi_env env, napi_status status, void* data) {
    job_t* j = (job_t*)data;
    napi_value v, msg;
    if (status == napi_ok && j->rc == 0) {
        napi_create_string_utf8(env, j->res.text, j->res.text_len, &v);
        napi_resolve_deferred(env, j->deferred, v);
        rt_result_free(&j->res);
    } else {
        napi_create_string_utf8(env, status == napi_ok ? j->err : "generation was cancelled", NAPI_AUTO_LENGTH, &msg);
        napi_create_error(env, NULL, msg, &v);
        napi_reject_deferred(env, j->deferred, v);
    }
    napi_delete_async_work(env, j->work);
    free(j->prompt);
    free(j);
}

static void engine_finalize(napi_env env, void* data, void* hint) {
    addon_engine* ae = (addon_engine*)data;
    (void)env;
    (void)hint;
    if (ae) {
        if (ae->engine) rt_engine_destroy(ae->engine);
        free(ae);
    }
}

/* create(model, device, maxCtx, maxNewTokens?) */
static napi_value addon_create(napi_env env, napi_callback_info info) {
    size_t argc = 4, n = 0;
    napi_value a[4], out;
    char model[128];
    int32_t device = 0, max_ctx = 16384, max_new = 0;
    rt_config cfg;
    addon_engine* ae;
    if (napi_get_cb_info(env, info, &argc, a, NULL, NULL) != napi_ok || argc < 3) return throw_text(env, "create(model, device, maxCtx[, maxNewTokens])");
    if (napi_get_value_string_utf8(env, a[0], model, sizeof model, &n) != napi_ok) return throw_text(env, "model must be a string");
    napi_get_value_int32(env, a[1], &device);
    napi_get_value_int32(env, a[2], &max_ctx);
    if (argc > 3) napi_get_value_int32(env, a[3], &max_new);
    memset(&cfg, 0, sizeof cfg);
    if (rt_arch_by_name(model, &cfg) != 0) return throw_text(env, rt_last_error(NULL));     /* "... not found ..." -> not_installed */
    cfg.device = device;
    cfg.tp_size = 1;
    cfg.max_batch = 1;                              /* the reference issues one execute() at a time (orchestrator.ts:361-442) */
    cfg.max_ctx = max_ctx;
    cfg.flags = RT_FLAG_PREFIX_CACHE;               /* a knight's next prompt extends its last one (src/utils/prompt.ts:60-77) */
    ae = (addon_engine*)calloc(1, sizeof *ae);
    if (!ae) return throw_text(env, "out of memory");
    ae->max_new_tokens = max_new;
    if (rt_engine_create(&cfg, &ae->engine) != 0) {
        free(ae);
        return throw_text(env, rt_last_error(NULL));
    }
    if (napi_create_external(env, ae, engine_finalize, NULL, &out) != napi_ok) {
        engine_finalize(env, ae, NULL);
        return throw_text(env, "napi_create_external failed");
    }
    return out;
}

/* generate(engine, prompt, timeoutMs) -> Promise<string> */
static napi_value addon_generate(napi_env env, napi_callback_info info) {
    size_t argc = 3;
    napi_value a[3], promise, name;
    double timeout = 0.0;
    job_t* j;
    if (napi_get_cb_info(env, info, &argc, a, NULL, NULL) != napi_ok || argc < 3) return throw_text(env, "generate(engine, prompt, timeoutMs)");
    j = (job_t*)calloc(1, sizeof *j);
    if (!j) return throw_text(env, "out of memory");
    if (napi_get_value_external(env, a[0], (void**)&j->ae) != napi_ok || !j->ae || !j->ae->engine) {
        free(j);
        return throw_text(env, "engine handle expected");
    }
    if (napi_get_value_string_utf8(env, a[1], NULL, 0, &j->len) != napi_ok) {
        free(j);
        return throw_text(env, "prompt must be a string");
    }
    j->prompt = (char*)malloc(j->len + 1);
    if (!j->prompt) {
        free(j);
        return throw_text(env, "out of memory");
    }
    napi_get_value_string_utf8(env, a[1], j->prompt, j->len + 1, &j->len);
    napi_get_value_double(env, a[2], &timeout);
    j->deadline_ms = timeout > 0.0 ? (uint64_t)timeout : 0;
    napi_create_promise(env, &j->deferred, &promise);
    napi_create_string_utf8(env, "rt_generate", NAPI_AUTO_LENGTH, &name);
    napi_create_async_work(env, NULL, name, job_run, job_done, j, &j->work);
    napi_queue_async_work(env, j->work);
    return promise;
}

static napi_value addon_context_length(napi_env env, napi_callback_info info) {
    size_t argc = 1;
    napi_value a[1], out;
    addon_engine* ae = NULL;
    rt_model_info mi;
    napi_get_cb_info(env, info, &argc, a, NULL, NULL);
    if (argc < 1 || napi_get_value_external(env, a[0], (void**)&ae) != napi_ok || !ae) return throw_text(env, "engine handle expected");
    if (rt_model_info_get(ae->engine, &mi) != 0) return throw_text(env, rt_last_error(ae->engine));
    napi_create_int32(env, mi.context_length, &out);
    return out;
}

static napi_value addon_destroy(napi_env env, napi_callback_info info) {
    size_t argc = 1;
    napi_value a[1], undef;
    addon_engine* ae = NULL;
    napi_get_cb_info(env, info, &argc, a, NULL, NULL);
    if (argc >= 1 && napi_get_value_external(env, a[0], (void**)&ae) == napi_ok && ae && ae->engine) {
        rt_engine_destroy(ae->engine);               /* the finalizer then only frees the wrapper */
        ae->engine = NULL;
    }
    napi_get_undefined(env, &undef);
    return undef;
}

static napi_value addon_init(napi_env env, napi_value exports) {
    napi_property_descriptor props[] = {
        {"create", NULL, addon_create, NULL, NULL, NULL, napi_default, NULL},
        {"generate", NULL, addon_generate, NULL, NULL, NULL, napi_default, NULL},
        {"contextLength", NULL, addon_context_length, NULL, NULL, NULL, napi_default, NULL},
        {"destroy", NULL, addon_destroy, NULL, NULL, NULL, napi_default, NULL},
    };
    napi_define_properties(env, exports, sizeof props / sizeof props[0], props);
    return exports;
}

NAPI_MODULE(NODE_GYP_MODULE_NAME, addon_init)
