/* Force-included prelude for compiling the UNMODIFIED reference with g++ (oracle/_ref only).
 * The reference uses MSVC-isms (std::tanf, std::sqrtf, ... e.g. src/integrators/bdpt.h:52,
 * src/core/math.h:124) and relies on transitive includes that libstdc++ does not provide. */
#pragma once
#ifdef __cplusplus
#include <cmath>
#include <cstring>
#include <cassert>
#include <mutex>
#include <thread>
#include <functional>
#include <algorithm>
#include <limits>
namespace std { using ::tanf; using ::sqrtf; using ::cosf; using ::sinf; using ::powf; }
#endif
