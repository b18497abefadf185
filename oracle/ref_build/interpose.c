/* TEST INFRASTRUCTURE ONLY. Linked into oracle/_ref/ref_cpu so the reference's serial path can run
 * on a box without a GPU: renv::Canvas and procedural::Perlin allocate with cudaMallocManaged even
 * in serial mode (/root/reference/src/rayenv/canvas.cu:7, src/procedural/perlin.cu:87-88,105-108).
 * These definitions take precedence over the shared libcudart's. */
#include <stdlib.h>
int cudaMallocManaged(void** p, size_t size, unsigned int flags) {
    (void)flags;
    *p = calloc(size ? size : 1, 1);
    return *p ? 0 : 2;
}
int cudaFree(void* p) {
    free(p);
    return 0;
}
