/*
 * ORACLE -- TEST INFRASTRUCTURE ONLY.  Not part of the product path.
 *
 * CPU restatement of simple-knn distCUDA2
 * (/root/reference/submodules/simple-knn/simple_knn.cu:131-183,185-221,
 *  spatial.cu:15-26): for every point, the mean of the three smallest squared
 * distances to OTHER points (self excluded by index, duplicates count as 0).
 * The reference's Morton ordering and box pruning (simple_knn.cu:45-129) only
 * accelerate the search -- the result is the exact 3-NN value -- so the oracle
 * is the brute-force definition.  dist = dx*dx + dy*dy + dz*dz is evaluated
 * with the FMA placement nvcc uses for updateKBest (simple_knn.cu:134-135):
 * t = dy*dy; t = fma(dx,dx,t); t = fma(dz,dz,t)  (read off the reference PTX).
 * "best" slots that are never filled stay FLT_MAX as in the reference
 * (P < 4), so the mean overflows to +inf exactly as the reference does.
 */
#include <float.h>
#include <math.h>

void ora_dist2_knn3(int P, const float* pts, float* out)
{
#pragma omp parallel for schedule(static)
    for (int i = 0; i < P; i++) {
        float best[3] = {FLT_MAX, FLT_MAX, FLT_MAX};
        const float px = pts[3 * i], py = pts[3 * i + 1], pz = pts[3 * i + 2];
        for (int j = 0; j < P; j++) {
            if (j == i) continue;
            float dx = pts[3 * j] - px, dy = pts[3 * j + 1] - py, dz = pts[3 * j + 2] - pz;
            float d = dy * dy; d = fmaf(dx, dx, d); d = fmaf(dz, dz, d);
            for (int k = 0; k < 3; k++)
                if (best[k] > d) { float t = best[k]; best[k] = d; d = t; }
        }
        out[i] = ((best[0] + best[1]) + best[2]) / 3.0f;
    }
}
