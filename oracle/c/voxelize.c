/* Oracle (test infrastructure only): sequential hard voxelisation, the public points_to_voxel algorithm of
 * SECOND / spconv 1.x VoxelGenerator / det3d VoxelGenerator. The reference calls it
 * through OPCD/pcdet/datasets/processor/data_processor.py:15-60,115-143 and clients/preprocess/voxelize.py:19-24,38-42;
 * the arithmetic itself lives in those un-vendored, un-pinned packages -> PARITY UNPINNED, restated here.
 * Build: gcc -O2 -ffp-contract=off -shared -fPIC (oracle/Makefile).
 */
#include <math.h>
#include <stdint.h>
#include <stdlib.h>
#include <string.h>

/* returns number of voxels M. coords are (z,y,x) like the generators return them (no batch column).
 * overflow_mode 0: `continue` when max_voxels is reached (spconv); 1: `break` (SECOND/det3d lineage).
 * f64_math 0: float arithmetic (points, range and voxel size are float32); 1: double arithmetic. */
int oracle_points_to_voxel(const float* pts, int n, int stride, int nfeat, const double* range, const double* vsize,
                           int max_pts, int max_vox, int overflow_mode, int f64_math,
                           float* voxels, int32_t* coords, int32_t* num) {
    int grid[3];
    for (int j = 0; j < 3; j++) {
        double g = f64_math ? (range[j + 3] - range[j]) / vsize[j]
                            : (double)(((float)range[j + 3] - (float)range[j]) / (float)vsize[j]);
        grid[j] = (int)nearbyint(g);
    }
    size_t cells = (size_t)grid[0] * grid[1] * grid[2];
    int32_t* cell_to_voxel = (int32_t*)malloc(cells * sizeof(int32_t));
    if (!cell_to_voxel) return -1;
    memset(cell_to_voxel, 0xff, cells * sizeof(int32_t));
    memset(voxels, 0, (size_t)max_vox * max_pts * nfeat * sizeof(float));
    memset(num, 0, (size_t)max_vox * sizeof(int32_t));
    int m = 0;
    for (int i = 0; i < n; i++) {
        const float* p = pts + (size_t)i * stride;
        int c[3], failed = 0;
        for (int j = 0; j < 3; j++) {
            double f;
            if (f64_math) f = floor(((double)p[j] - range[j]) / vsize[j]);
            else f = (double)floorf((p[j] - (float)range[j]) / (float)vsize[j]);
            if (!(f >= 0 && f < grid[j])) { failed = 1; break; }
            c[j] = (int)f;
        }
        if (failed) continue;
        size_t key = ((size_t)c[2] * grid[1] + c[1]) * grid[0] + c[0];
        int v = cell_to_voxel[key];
        if (v == -1) {
            if (m >= max_vox) { if (overflow_mode) break; else continue; }
            v = m++;
            cell_to_voxel[key] = v;
            coords[v * 3 + 0] = c[2]; coords[v * 3 + 1] = c[1]; coords[v * 3 + 2] = c[0];
        }
        if (num[v] < max_pts) {
            memcpy(voxels + ((size_t)v * max_pts + num[v]) * nfeat, p, nfeat * sizeof(float));
            num[v]++;
        }
    }
    free(cell_to_voxel);
    return m;
}
