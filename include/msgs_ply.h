/* msgs_ply.h — C ABI of libmsgs_ply.so: host-side loader of MS-GS model files (the on-disk format either
 * side of the rasterizer path: the viewer / render.py path reads a trained model and renders it forward-only).
 *
 * Replaces GaussianModel.load_ply (/root/reference/scene/gaussian_model.py:358-417; plyfile + ~15 float64 numpy
 * temporaries + torch.tensor(...).cuda() per attribute) and the per-call torch.cat of get_features (:127-153) for that
 * path: the binary little-endian PLY written by GaussianModel.save_ply (:314-344) is memory-mapped and converted in one
 * multi-threaded pass into the fp32 layout the rasterizer takes, in caller-owned (ideally pinned) buffers.
 * Pure host code (no CUDA).  All functions are thread-safe per handle; errors: NULL / negative code + msgs_ply_last_error(). */
#ifndef MSGS_PLY_H
#define MSGS_PLY_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define MSGS_PLY_OK 0
#define MSGS_PLY_ERR_ARG (-1)     /* NULL handle / required buffer */
#define MSGS_PLY_ERR_IO (-2)      /* open / stat / mmap failed, truncated file */
#define MSGS_PLY_ERR_FORMAT (-3)  /* not a binary_little_endian PLY with the vertex attributes named above */

typedef struct msgs_ply_info {
    int64_t num_points;       /* P */
    int sh_coeffs;            /* M = 1 + (number of f_rest_*) / 3 (16 for max_sh_degree 3) */
    int n_lvl_occ;            /* L: number of occ_multiplier_* columns (0 in a plain 3DGS file) */
    int n_lvl_dc;             /* D: number of dc_delta_*_* columns / 3 */
    int has_ms_attributes;    /* base_gaussian_mask, max_pixel_sizes and min_pixel_sizes all present */
    int64_t row_bytes;        /* bytes per vertex in the file */
    int64_t data_bytes;       /* P * row_bytes */
} msgs_ply_info;

/* Memory-maps `path`, parses the header, fills *info.  Returns a handle or NULL. */
void* msgs_ply_open(const char* path, msgs_ply_info* info);

/* Converts all vertices.  Buffer shapes (fp32 unless noted; any pointer may be NULL = not wanted):
 *   means3D (P,3) | shs (P,M,3): shs[p][0][c] = f_dc_c, shs[p][k][c] = f_rest_{c*(M-1)+k-1} — the (P,M,3) tensor
 *   get_features returns | opacity (P) raw | scales (P,3) raw | rotations (P,4) raw | occ_multiplier (P,L) |
 *   dc_delta (P,3D) | base_mask (P) uint8 | max_pixel_sizes (P) | min_pixel_sizes (P).
 * For a file without MS-GS attributes occ_multiplier / dc_delta are left untouched (L = D = 0) and base_mask = 0,
 * max = min = -1.  `threads` <= 0: hardware concurrency (at most 16). */
int msgs_ply_read(void* handle, float* means3D, float* shs, float* opacity, float* scales, float* rotations,
                  float* occ_multiplier, float* dc_delta, uint8_t* base_mask, float* max_pixel_sizes,
                  float* min_pixel_sizes, int threads);

void msgs_ply_close(void* handle);
const char* msgs_ply_last_error(void);

#ifdef __cplusplus
}
#endif
#endif /* MSGS_PLY_H */
