/* The auxiliary-library functions bindings/lua_mrgpu.c uses (Lua 5.2 reference manual, section 5.1). */
#ifndef LAUXLIB_STUB_H
#define LAUXLIB_STUB_H
#include "lua.h"
typedef struct luaL_Reg { const char *name; lua_CFunction func; } luaL_Reg;
void *luaL_checkudata(lua_State *L, int arg, const char *tname);
int luaL_error(lua_State *L, const char *fmt, ...);
const char *luaL_checklstring(lua_State *L, int arg, size_t *l);
lua_Integer luaL_checkinteger(lua_State *L, int arg);
void luaL_checktype(lua_State *L, int arg, int t);
void luaL_setmetatable(lua_State *L, const char *tname);
int luaL_newmetatable(lua_State *L, const char *tname);
void luaL_setfuncs(lua_State *L, const luaL_Reg *l, int nup);
#endif
