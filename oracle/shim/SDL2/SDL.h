/* Build-dir shim: the reference includes "SDL2/SDL.h"; its vendored SDL2 tree has SDL.h at
 * externals/opengl/SDL2/include/SDL.h. Headers only - nothing of SDL is ever executed on the
 * offline (BDPT) path. Test infrastructure; not shipped. */
#pragma once
#include <SDL.h>
