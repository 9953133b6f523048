/* POD vector types of the C API (reference include/c/vkt/linalg.h:11-80, hot-path subset). */
#ifndef VKT_LINALG_H
#define VKT_LINALG_H

typedef struct { float x, y; } vktVec2f_t;
typedef struct { float x, y, z; } vktVec3f_t;
typedef struct { int x, y, z; } vktVec3i_t;
typedef struct { vktVec3f_t min, max; } vktBox3f_t;

typedef enum { vktAxisX, vktAxisY, vktAxisZ } vktAxis;   /* reference include/c/vkt/linalg.h:94-99 */

#endif
