/* Declarations of exactly the Lua 5.2 C-API functions bindings/lua_mrgpu.c uses, with the signatures
 * of the Lua 5.2 reference manual (section 4.8), so that the binding can be syntax- and type-checked
 * in an image that has no Lua.  Not a Lua header: nothing here is implemented. */
#ifndef LUA_STUB_H
#define LUA_STUB_H
#include <stddef.h>
typedef struct lua_State lua_State;
typedef double lua_Number;
typedef ptrdiff_t lua_Integer;
typedef int (*lua_CFunction)(lua_State *L);
#define LUA_TNIL 0
#define LUA_TTABLE 5
#define LUA_REGISTRYINDEX (-1000000 - 1000)
#define lua_upvalueindex(i) (LUA_REGISTRYINDEX - (i))
int lua_type(lua_State *L, int idx);
#define lua_isnil(L, n) (lua_type(L, (n)) == LUA_TNIL)
void lua_getfield(lua_State *L, int idx, const char *k);
void lua_setfield(lua_State *L, int idx, const char *k);
void lua_settop(lua_State *L, int idx);
int lua_toboolean(lua_State *L, int idx);
void *lua_newuserdata(lua_State *L, size_t sz);
void *lua_touserdata(lua_State *L, int idx);
void lua_createtable(lua_State *L, int narr, int nrec);
void lua_pushnumber(lua_State *L, lua_Number n);
void lua_pushinteger(lua_State *L, lua_Integer n);
const char *lua_pushlstring(lua_State *L, const char *s, size_t len);
void lua_pushnil(lua_State *L);
void lua_pushcclosure(lua_State *L, lua_CFunction fn, int n);
void lua_pushvalue(lua_State *L, int idx);
void lua_rawseti(lua_State *L, int idx, int n);
#endif
