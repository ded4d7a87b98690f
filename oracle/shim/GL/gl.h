/* Build-dir shim: empty <GL/gl.h>. GLEW (vendored by the reference) already declares what the
 * reference's realtime passes name; none of it runs on the offline (BDPT) path. */
#pragma once
