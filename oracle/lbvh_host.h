/* lbvh_host.h — CPU ORACLE, TEST INFRASTRUCTURE ONLY: sequential host build of the LBVH / BVH8
 * (see lbvh_host.c). Only tests/ may load this. */
#ifndef LBVH_HOST_H
#define LBVH_HOST_H
#include <stdint.h>
#include "rtb200.h"
#ifdef __cplusplus
extern "C" {
#endif
typedef struct OrBvh OrBvh;
OrBvh* or_bvh_build(const RtSceneDesc* desc);
void or_bvh_free(OrBvh* b);
int32_t or_bvh_n_prims(const OrBvh* b);
int32_t or_bvh_n_nodes8(const OrBvh* b);
int32_t or_bvh_levels(const OrBvh* b);
const uint64_t* or_bvh_morton(const OrBvh* b);      /* n, sorted */
const int32_t* or_bvh_perm(const OrBvh* b);         /* n: sorted position -> prim (instance-major index) */
const int32_t* or_bvh_left(const OrBvh* b);         /* n-1 */
const int32_t* or_bvh_right(const OrBvh* b);        /* n-1 */
const int32_t* or_bvh_parent(const OrBvh* b);       /* 2n-1 */
const float* or_bvh_node_boxes(const OrBvh* b);     /* 6*(2n-1) */
const float* or_bvh_bounds(const OrBvh* b);         /* 6 */
const uint32_t* or_bvh_nodes8(const OrBvh* b);      /* 20 words (80 bytes) per node */
const float* or_bvh_prims8(const OrBvh* b);         /* 12 floats (48 bytes) per triangle, leaf order */
const int32_t* or_bvh_prim_order8(const OrBvh* b);  /* n */
#ifdef __cplusplus
}
#endif
#endif
