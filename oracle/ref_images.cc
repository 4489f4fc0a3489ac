// The reference's own image decoders compiled for the host from where they lie under /root/reference: stb_image
// (dependencies/stb_image/stb_image.h, what load_stbi / stbi_load hands the loader, src/common_host.cu:214-217 and
// nerf_loader.cu:571) and tinyexr (dependencies/tinyexr/tinyexr.h with its bundled miniz, what load_exr /
// load_exr_to_gpu read, src/tinyexr_wrapper.cu:125-260). TEST INFRASTRUCTURE: tests/test_host_io_cpu.py decodes the same
// files with csrc/host_png.cu / host_image.cu / host_exr.cu and compares texel for texel; nothing of the product links it.
// Built by oracle/build_ref.sh into oracle/_ref/libref_images.so.
#define STB_IMAGE_IMPLEMENTATION
#include <stb_image.h>

#define TINYEXR_IMPLEMENTATION
#include <tinyexr.h>

#include <cstdint>
#include <cstring>

// stbi_load(path, &w, &h, &comp, 4): RGBA8 rows top to bottom. Returns 0, or 1 with the failure reason in `why`.
extern "C" int ref_stbi_load(const char* path, int* w, int* h, int* comp, uint8_t* out, size_t cap, char* why, size_t why_cap) {
	uint8_t* img = stbi_load(path, w, h, comp, 4);
	if (!img) {
		if (why && why_cap) { strncpy(why, stbi_failure_reason(), why_cap - 1); why[why_cap - 1] = 0; }
		return 1;
	}
	const size_t bytes = (size_t)*w * *h * 4;
	if (out && cap >= bytes) memcpy(out, img, bytes);
	stbi_image_free(img);
	return 0;
}

// stbi_loadf(path, &w, &h, &comp, 4): float RGBA (Radiance HDR files as they are, others through stb_image's ldr-to-hdr curve).
extern "C" int ref_stbi_loadf(const char* path, int* w, int* h, float* out, size_t cap_floats, char* why, size_t why_cap) {
	int comp = 0;
	float* img = stbi_loadf(path, w, h, &comp, 4);
	if (!img) {
		if (why && why_cap) { strncpy(why, stbi_failure_reason(), why_cap - 1); why[why_cap - 1] = 0; }
		return 1;
	}
	const size_t n = (size_t)*w * *h * 4;
	if (out && cap_floats >= n) memcpy(out, img, n * sizeof(float));
	stbi_image_free(img);
	return 0;
}

// LoadEXR (the simple API behind load_exr): float RGBA rows top to bottom.
extern "C" int ref_tinyexr_load(const char* path, int* w, int* h, float* out, size_t cap_floats, char* why, size_t why_cap) {
	float* rgba = nullptr;
	const char* err = nullptr;
	if (LoadEXR(&rgba, w, h, path, &err) != TINYEXR_SUCCESS) {
		if (why && why_cap) { strncpy(why, err ? err : "?", why_cap - 1); why[why_cap - 1] = 0; }
		if (err) FreeEXRErrorMessage(err);
		return 1;
	}
	const size_t n = (size_t)*w * *h * 4;
	if (out && cap_floats >= n) memcpy(out, rgba, n * sizeof(float));
	free(rgba);
	return 0;
}
