/* Oracle (TEST INFRASTRUCTURE / CPU baseline): plain-C restatement of the reference's projection stage.
 *
 * Follows reference ts2d/core/util/image.py:75-93 (sitk.MaximumProjectionImageFilter and
 * sitk.MeanProjectionImageFilter with SetProjectionDimension(1), one filter execution per channel, i.e. the
 * volume is streamed twice) and ts2d/tool.py:184 (Cast to float32).  ITK's accumulators: running max in the
 * pixel type; sum in double divided by the line length.  ITK threads its filters over output regions; the
 * OpenMP loop over z mirrors that so the timed CPU baseline uses every host core like the reference would.
 * Input int16 [nz][ny][nx] already in RAI order; output float32 [nz][nx].
 */
#include <stdint.h>
#include <stddef.h>

void oracle_project_max_i16(const int16_t *vol, int64_t nz, int64_t ny, int64_t nx, float *out) {
#pragma omp parallel for schedule(static)
    for (int64_t z = 0; z < nz; ++z) {
        const int16_t *plane = vol + z * ny * nx;
        for (int64_t x = 0; x < nx; ++x) {
            int16_t m = plane[x];
            for (int64_t y = 1; y < ny; ++y) {
                const int16_t v = plane[y * nx + x];
                if (v > m) m = v;
            }
            out[z * nx + x] = (float)m;
        }
    }
}

void oracle_project_mean_i16(const int16_t *vol, int64_t nz, int64_t ny, int64_t nx, float *out) {
#pragma omp parallel for schedule(static)
    for (int64_t z = 0; z < nz; ++z) {
        const int16_t *plane = vol + z * ny * nx;
        for (int64_t x = 0; x < nx; ++x) {
            double s = 0.0;
            for (int64_t y = 0; y < ny; ++y) s += (double)plane[y * nx + x];
            out[z * nx + x] = (float)(s / (double)ny);
        }
    }
}
