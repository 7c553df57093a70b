/*
 * mobilert_b200_test.h — C-ABI of libmobilert_b200_test.so: the primitive-level device functions of the hot path, one call
 * at a time, so that the reference's own known-answer vectors can be run against the device code. A separate library on
 * purpose: nothing of this is in libmobilert_b200.so, which a front-end links.
 */
#ifndef MOBILERT_B200_TEST_H
#define MOBILERT_B200_TEST_H

#include "mobilert_b200.h"

#ifdef __cplusplus
extern "C" {
#endif

/* Primitive-level device functions exposed for the reference's known-answer tests (Unit_Testing/TestTriangle.cpp,
 * TestAABB.cpp, TestPlane.cpp). Each runs the device function in a one-thread kernel. */
int mrt_test_triangle_intersect(const mrt_triangle *tri, const float *o, const float *d, float t_max, int32_t self_skip,
                                float *out_t, float *out_normal_uv_point8);
int mrt_test_sphere_intersect(const mrt_sphere *s, const float *o, const float *d, float t_max, float *out_t, float *out_normal);
int mrt_test_plane_intersect(const mrt_plane *p, const float *o, const float *d, float t_max, float *out_t);
int mrt_test_aabb_intersect(const float *box_min, const float *box_max, const float *o, const float *d, int32_t *out_hit);
int mrt_test_incremental_avg(const float *rgb, int32_t avg, int32_t num_sample, int32_t *out);
/* MobileRT::haltonSequence (MobileRT/Utils/Utils.cpp:43-53) evaluated on the device for indices [first, first+n). */
int mrt_test_halton(uint32_t first, int64_t n, uint32_t base, float *out);

/* RegularGrid<T>::RegularGrid + addPrimitives (MobileRT/Accelerators/RegularGrid.hpp:112-283, with T::intersect(AABB) of
 * Plane.cpp:146-155, Sphere.cpp:102-123, Triangle.cpp:142-229) as libmobilert_b200.so builds it on the host for
 * MRT_ACC_REGULAR_GRID. kind: 0 planes, 1 spheres, 2 triangles. out_params12: bounds min, bounds max, cellSizeInverted_,
 * cellSize_. out_start: 32^3 + 1 offsets; out_items (may be null): primitive indices cell after cell, ascending inside a
 * cell, at most items_cap of them; *out_n_items: how many there are. Runs on the host alone (no device needed). */
int mrt_test_build_grid(const mrt_scene_desc *scene, int32_t kind, float *out_params12, uint32_t *out_start, int32_t *out_items, int64_t items_cap,
                        int64_t *out_n_items);

/* Description of the last failure of an entry point of this library on the calling thread. */
const char *mrt_test_last_error(void);

#ifdef __cplusplus
}
#endif
#endif /* MOBILERT_B200_TEST_H */
