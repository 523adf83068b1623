/* lua_mrgpu.c -- Lua 5.2 C-API module over include/mrgpu.h: the glue a maintainer of lua-mapreduce
 * would add next to mapreduce/job.lua (SURVEY f-1).  Not built here (this image has no lua.h); the
 * test suite only checks that it compiles against the API it names (tests/lua_stub declares
 * exactly the Lua C-API functions used).  Build where Lua 5.2 is installed:
 *     gcc -O2 -fPIC -shared -I include bindings/lua_mrgpu.c -o mrgpu.so -L lua-mapreduce-1_b200 -lmrgpu -llua5.2
 *
 *   local mrgpu = require "mrgpu"
 *   local ctx = mrgpu.new{ num_reducers = 15, arena_bytes = 2^28 }       -- one per worker process and GPU
 *   ctx:job_begin(false)                                                 -- worker.lua:62-63
 *   ctx:map_file(map_key, path)    -- or ctx:map_string(map_key, s)      -- job.lua:166-227
 *   local info = ctx:finish()                                            -- emit+combine+partition+sort+reduce
 *   for key, values in ctx:pairs(part) do ... end                        -- server.lua:358-383, values = {count}
 *
 * Every failing call raises a Lua error carrying mr_last_error, so the reference's xpcall ->
 * BROKEN -> retry machinery (worker.lua:116-137) sees it like any failing job.
 */
#include <string.h>

#include "lua.h"
#include "lauxlib.h"

#include "mrgpu.h"

#define CTX_MT "mrgpu.ctx"
#define RES_MT "mrgpu.result"

typedef struct { mr_ctx *ctx; } ctx_box;
typedef struct { mr_result *res; } res_box;

static mr_ctx *check_ctx(lua_State *L)
{
    ctx_box *b = (ctx_box *)luaL_checkudata(L, 1, CTX_MT);
    if (!b->ctx) luaL_error(L, "mrgpu: context is closed");
    return b->ctx;
}

static void check_rc(lua_State *L, mr_ctx *ctx, int rc, const char *what)
{
    if (rc != MR_OK) luaL_error(L, "%s: %s", what, mr_last_error(ctx));
}

static uint32_t opt_kind(lua_State *L, int t, const char *field, const char *const names[], uint32_t dflt)
{
    uint32_t v = dflt;
    lua_getfield(L, t, field);
    if (!lua_isnil(L, -1)) {
        const char *s = luaL_checklstring(L, -1, NULL);
        int found = 0;
        for (uint32_t i = 0; names[i]; ++i)
            if (strcmp(names[i], s) == 0) { v = i; found = 1; }
        if (!found) luaL_error(L, "mrgpu.new: unknown %s kind '%s'", field, s);
    }
    lua_settop(L, -2);
    return v;
}

static lua_Integer opt_int_field(lua_State *L, int t, const char *field, lua_Integer dflt)
{
    lua_getfield(L, t, field);
    lua_Integer v = lua_isnil(L, -1) ? dflt : luaL_checkinteger(L, -1);
    lua_settop(L, -2);
    return v;
}

/* mrgpu.new{ num_reducers=, arena_bytes=, max_keys=, device=, world_size=, rank=,
 *            map = "tokenize_ws"|"records_u32"|"bigram_ws", reduce = "sum_u64"|"min_u64"|"max_u64"|"count",
 *            partition = "fnv1_lua52_double"|"mod_u32" }   -- the task module's `native` descriptor */
static int l_new(lua_State *L)
{
    static const char *const maps[] = { "tokenize_ws", "records_u32", "bigram_ws", NULL };
    static const char *const reds[] = { "sum_u64", "min_u64", "max_u64", "count", NULL };
    static const char *const parts[] = { "fnv1_lua52_double", "mod_u32", NULL };
    luaL_checktype(L, 1, LUA_TTABLE);
    mr_config cfg;
    memset(&cfg, 0, sizeof cfg);
    cfg.struct_size = sizeof cfg;
    cfg.device = (int32_t)opt_int_field(L, 1, "device", 0);
    cfg.num_reducers = (uint32_t)opt_int_field(L, 1, "num_reducers", 10);
    cfg.map_kind = opt_kind(L, 1, "map", maps, MR_MAP_TOKENIZE_WS);
    cfg.reduce_kind = opt_kind(L, 1, "reduce", reds, MR_REDUCE_SUM_U64);
    cfg.partition_kind = opt_kind(L, 1, "partition", parts, MR_PARTITION_FNV1_LUA52);
    cfg.arena_bytes = (uint64_t)opt_int_field(L, 1, "arena_bytes", 1 << 28);
    cfg.max_keys = (uint64_t)opt_int_field(L, 1, "max_keys", 0);
    cfg.world_size = (uint32_t)opt_int_field(L, 1, "world_size", 1);
    cfg.rank = (uint32_t)opt_int_field(L, 1, "rank", 0);
    ctx_box *b = (ctx_box *)lua_newuserdata(L, sizeof *b);
    b->ctx = NULL;
    luaL_setmetatable(L, CTX_MT);
    if (mr_ctx_create(&cfg, &b->ctx) != MR_OK) luaL_error(L, "mrgpu.new: %s", mr_last_error(NULL));
    return 1;
}

static int l_close(lua_State *L)
{
    ctx_box *b = (ctx_box *)luaL_checkudata(L, 1, CTX_MT);
    if (b->ctx) { mr_ctx_destroy(b->ctx); b->ctx = NULL; }
    return 0;
}

static int l_job_begin(lua_State *L)
{
    mr_ctx *ctx = check_ctx(L);
    check_rc(L, ctx, mr_job_begin(ctx, lua_toboolean(L, 2)), "job_begin");
    return 0;
}

static int l_map_string(lua_State *L) /* ctx:map_string(map_key, bytes) */
{
    mr_ctx *ctx = check_ctx(L);
    size_t n = 0;
    const char *s = luaL_checklstring(L, 3, &n);
    check_rc(L, ctx, mr_map_split(ctx, (uint64_t)luaL_checkinteger(L, 2), s, (uint64_t)n), "map_string");
    return 0;
}

static int l_map_file(lua_State *L) /* ctx:map_file(map_key, path): what mapfn.lua:4 does with io.lines(value) */
{
    mr_ctx *ctx = check_ctx(L);
    check_rc(L, ctx, mr_map_split_path(ctx, (uint64_t)luaL_checkinteger(L, 2), luaL_checklstring(L, 3, NULL)), "map_file");
    return 0;
}

static void push_info(lua_State *L, const mr_job_info *info, const mr_stats *st)
{
    lua_createtable(L, 0, 8);
    lua_pushnumber(L, (lua_Number)info->n_keys);      lua_setfield(L, -2, "n_keys");
    lua_pushnumber(L, (lua_Number)info->n_tokens);    lua_setfield(L, -2, "n_tokens");
    lua_pushnumber(L, (lua_Number)info->input_bytes); lua_setfield(L, -2, "input_bytes");
    lua_pushnumber(L, (lua_Number)info->map_keys);    lua_setfield(L, -2, "map_keys");
    lua_pushnumber(L, (lua_Number)info->tie_rounds);  lua_setfield(L, -2, "tie_rounds");
    /* what job_mark_as_written stores as cpu_time / real_time (job.lua:134-152), in seconds */
    lua_pushnumber(L, (lua_Number)st->map_ms * 1e-3);    lua_setfield(L, -2, "map_time");
    lua_pushnumber(L, (lua_Number)st->reduce_ms * 1e-3); lua_setfield(L, -2, "reduce_time");
}

static int l_finish(lua_State *L) /* ctx:finish() -> info table */
{
    mr_ctx *ctx = check_ctx(L);
    mr_job_info info;
    mr_stats st;
    check_rc(L, ctx, mr_map_finish(ctx), "map_finish");
    check_rc(L, ctx, mr_reduce_all(ctx), "reduce_all");
    check_rc(L, ctx, mr_job_sync(ctx, &info), "job_sync");
    check_rc(L, ctx, mr_stats_get(ctx, &st), "stats");
    push_info(L, &info, &st);
    return 1;
}

static int l_res_gc(lua_State *L)
{
    res_box *r = (res_box *)luaL_checkudata(L, 1, RES_MT);
    if (r->res) { mr_result_release(r->res); r->res = NULL; }
    return 0;
}

static int pairs_next(lua_State *L) /* upvalue 1: the result cursor */
{
    res_box *r = (res_box *)lua_touserdata(L, lua_upvalueindex(1));
    const char *key;
    uint64_t klen, count;
    if (!r->res || mr_result_next(r->res, &key, &klen, &count) != MR_OK) {
        if (r->res) { mr_result_release(r->res); r->res = NULL; }
        lua_pushnil(L);
        return 1;
    }
    lua_pushlstring(L, key, (size_t)klen);
    lua_createtable(L, 1, 0);                 /* values: the reference hands finalfn a table {count} */
    lua_pushnumber(L, (lua_Number)count);
    lua_rawseti(L, -2, 1);
    return 2;
}

static int l_pairs(lua_State *L) /* for key, values in ctx:pairs(part) do ... end */
{
    mr_ctx *ctx = check_ctx(L);
    res_box *r = (res_box *)lua_newuserdata(L, sizeof *r);
    r->res = NULL;
    luaL_setmetatable(L, RES_MT);
    check_rc(L, ctx, mr_reduce_partition(ctx, (uint32_t)luaL_checkinteger(L, 2), &r->res), "pairs");
    lua_pushcclosure(L, pairs_next, 1);
    return 1;
}

static int l_partition_of(lua_State *L) /* mrgpu.partition_of(key, num_reducers): partitionfn.lua:8-16 */
{
    size_t n = 0;
    const char *s = luaL_checklstring(L, 1, &n);
    lua_pushinteger(L, (lua_Integer)mr_partition_of(s, (uint64_t)n, (uint32_t)luaL_checkinteger(L, 2)));
    return 1;
}

static const luaL_Reg ctx_methods[] = {
    { "job_begin", l_job_begin }, { "map_string", l_map_string }, { "map_file", l_map_file },
    { "finish", l_finish }, { "pairs", l_pairs }, { "close", l_close }, { "__gc", l_close }, { NULL, NULL }
};

static const luaL_Reg module_functions[] = {
    { "new", l_new }, { "partition_of", l_partition_of }, { NULL, NULL }
};

int luaopen_mrgpu(lua_State *L)
{
    luaL_newmetatable(L, CTX_MT);
    lua_pushvalue(L, -1);
    lua_setfield(L, -2, "__index");
    luaL_setfuncs(L, ctx_methods, 0);
    luaL_newmetatable(L, RES_MT);
    lua_pushcclosure(L, l_res_gc, 0);
    lua_setfield(L, -2, "__gc");
    lua_createtable(L, 0, 2);
    luaL_setfuncs(L, module_functions, 0);
    return 1;
}
