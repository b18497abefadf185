// Build shim (test infrastructure): the reference only needs the SDL_Surface type name and
// SDL_CreateRGBSurfaceFrom's signature (include/rayenv/canvas.h:11,33; src/rayenv/canvas.cu:23-29).
// No window / graphics stack exists on this box; the surface call is never reached.
#ifndef ORACLE_SHIM_SDL_H
#define ORACLE_SHIM_SDL_H
#include <cstdint>
struct SDL_Surface { int unused; };
typedef uint32_t Uint32;
inline SDL_Surface* SDL_CreateRGBSurfaceFrom(void*, int, int, int, int, Uint32, Uint32, Uint32, Uint32) { return nullptr; }
#endif
