// Output stage of the reference's main(): kernel.cu:763-778.
// Reinhard c/(c+1) (color.h:59-62), gamma with the exponent (float)(1/2.2) evaluated by double pow
// (color.h:68-71, kernel.cu:774), (int)(c*255) truncation, ASCII P3 on one line, x mirrored (kernel.cu:768).
// The accumulator this library hands out is row-major, so the Morton lookup of kernel.cu:771 disappears.
#include "host_common.h"

using namespace cupt;

static inline void tonemap_px(const float* rgb, int32_t* out)
{
	const float a = (float)(1 / 2.2);
	for (int k = 0; k < 3; ++k) {
		double c = (double)rgb[k];
		c = c / (c + 1);
		c = pow(c, (double)a);
		out[k] = (int32_t)(c * 255);
	}
}

extern "C" {

int cupt_tonemap(const float* rgb_mean, int64_t npixels, int32_t* out_rgb255)
{
	if (!rgb_mean || !out_rgb255 || npixels < 0)
		return set_error(CUPT_ERR_INVALID, "bad arguments");
	for (int64_t i = 0; i < npixels; ++i)
		tonemap_px(rgb_mean + 3 * i, out_rgb255 + 3 * i);
	return CUPT_OK;
}

int cupt_write_ppm(const char* path, const float* rgb_mean, int width, int height)
{
	if (!path || !rgb_mean || width <= 0 || height <= 0)
		return set_error(CUPT_ERR_INVALID, "bad arguments");
	FILE* fp = fopen(path, "w");
	if (!fp)
		return set_error(CUPT_ERR_IO, "cannot open %s for writing", path);
	fprintf(fp, "P3 %d %d 255\n", width, height);
	for (int y = 0; y < height; ++y)
		for (int x = width - 1; x >= 0; --x) {
			int32_t c[3];
			tonemap_px(rgb_mean + 3 * ((size_t)y * width + x), c);
			fprintf(fp, "%d %d %d ", c[0], c[1], c[2]);
		}
	int rc = fclose(fp);
	return rc == 0 ? CUPT_OK : set_error(CUPT_ERR_IO, "write to %s failed", path);
}

int cupt_write_ppm8(const char* path, const uint8_t* rgb, int width, int height)
{
	if (!path || !rgb || width <= 0 || height <= 0)
		return set_error(CUPT_ERR_INVALID, "bad arguments");
	FILE* fp = fopen(path, "w");
	if (!fp)
		return set_error(CUPT_ERR_IO, "cannot open %s for writing", path);
	fprintf(fp, "P3 %d %d 255\n", width, height);
	const size_t n = (size_t)width * height;
	for (size_t i = 0; i < n; ++i) fprintf(fp, "%d %d %d ", rgb[3 * i], rgb[3 * i + 1], rgb[3 * i + 2]);
	int rc = fclose(fp);
	return rc == 0 ? CUPT_OK : set_error(CUPT_ERR_IO, "write to %s failed", path);
}

// The raw accumulator as a Portable Float Map ("PF", three f32 per pixel, little-endian = negative scale, rows bottom to top by
// the format's convention): no tonemap, no mirroring, every bit of the image survives the file.  SURVEY 8(f) rank 4; the
// reference itself only writes the tonemapped P3 text (kernel.cu:763-778).
int cupt_write_pfm(const char* path, const float* rgb, int width, int height)
{
	if (!path || !rgb || width <= 0 || height <= 0)
		return set_error(CUPT_ERR_INVALID, "bad arguments");
	FILE* fp = fopen(path, "wb");
	if (!fp)
		return set_error(CUPT_ERR_IO, "cannot open %s for writing", path);
	fprintf(fp, "PF\n%d %d\n-1.0\n", width, height);
	size_t written = 0;
	for (int y = height - 1; y >= 0; --y)
		written += fwrite(rgb + 3 * (size_t)y * width, sizeof(float), 3 * (size_t)width, fp);
	int rc = fclose(fp);
	return (rc == 0 && written == 3 * (size_t)width * height) ? CUPT_OK : set_error(CUPT_ERR_IO, "write to %s failed", path);
}

}  // extern "C"
