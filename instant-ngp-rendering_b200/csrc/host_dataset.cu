// Testbed::load_training_data for NeRF scenes in host C++ (src/testbed.cu:125-152 -> load_nerf, src/nerf_loader.cu:273-740:
// frame ordering and culling :331-396, read_focal_length :243-271, read_lens :175-241, scale / offset / aabb :448-506,
// nerf_matrix_to_ngp nerf_loader.h:101-120, the byte-image fix-ups of convert_rgba32 :41-63 and :576-618) for the lens
// models the NeRF path implements (Perspective and OpenCV).
//
// B200 side of it: JPEG files - the fox scene is 50 of them - are decoded ON THE GPU by nvjpeg straight into device
// memory and expanded to the RGBA8 training image by a kernel, so the pixels never exist on the host; PNG files are
// inflated on the host (host_png.cu) and uploaded, EXR files (HDR scenes, the image mode's albert.exr) are read by
// host_exr.cu. An image file instead of a scene selects the neural-image mode, as Testbed::load_file does.
// The Python mirror (loader.py) is a second implementation of the same rules; tests compare the two.
#include "testbed.h"
#include "host_io.h"

#include <nvjpeg.h>

#include <algorithm>
#include <cmath>
#include <cstring>
#include <dirent.h>

namespace ngpb {

// SI::natural::compare: digit runs compare by value, the rest by character
// SI::natural::compare (dependencies/NaturalSort/natural_sort.hpp:170-236), the order of the reference's frames
// (src/nerf_loader.cu:347-349), restated with its quirks: characters compare case-folded as unsigned bytes; a digit run meets a
// digit run by value (leading zeros skipped) unless it follows a '.', where runs compare as fractions; a fraction against an
// integer run counts as "greater" and the reverse as "equal" (compare_number returns bool from an int function, :140-143);
// repeated spaces collapse. Pinned to the reference's own header by tests/golden/ref_natural_sort.json.gz.
static int natural_fold(char c) {
	const unsigned char u = (unsigned char)c;
	return u >= 'A' && u <= 'Z' ? u + 32 : u;
}
static bool natural_digit(char c) { return c >= '0' && c <= '9'; }
static int natural_number(const std::string& a, size_t l, size_t le, bool fa, const std::string& b, size_t r, size_t re, bool fb) {
	if (fa && !fb) return 1;
	if (!fa && fb) return 0;
	if (fa) {
		for (; l < le && r < re; ++l, ++r) if (a[l] != b[r]) return a[l] < b[r] ? -1 : 1;
		while (l < le && a[l] == '0') ++l;
		while (r < re && b[r] == '0') ++r;
		if (l == le && r != re) return -1;
		if (l != le && r == re) return 1;
		return 0;
	}
	while (l < le && a[l] == '0') ++l;
	while (r < re && b[r] == '0') ++r;
	if (le - l != re - r) return le - l < re - r ? -1 : 1;
	for (; l < le; ++l, ++r) if (a[l] != b[r]) return a[l] < b[r] ? -1 : 1;
	return 0;
}
static bool natural_less(const std::string& a, const std::string& b) {
	const size_t na = a.size(), nb = b.size();
	auto at = [](const std::string& s, size_t i) { return i < s.size() ? s[i] : '\0'; };
	size_t i = 0, j = 0;
	bool space_a = false, space_b = false;
	while (i < na && j < nb) {
		while (space_a && i < na && a[i] == ' ') ++i;
		space_a = at(a, i) == ' ';
		while (space_b && j < nb && b[j] == ' ') ++j;
		space_b = at(b, j) == ' ';
		const char ca = at(a, i), cb = at(b, j);
		if (!natural_digit(ca) || !natural_digit(cb)) {
			if (natural_fold(ca) != natural_fold(cb)) return natural_fold(ca) < natural_fold(cb);
			++i; ++j;
		} else {
			size_t ei = i, ej = j;
			while (ei < na && natural_digit(a[ei])) ++ei;
			while (ej < nb && natural_digit(b[ej])) ++ej;
			const int order = natural_number(a, i, ei, i > 0 && a[i - 1] == '.', b, j, ej, j > 0 && b[j - 1] == '.');
			if (order != 0) return order < 0;
			i = ei; j = ej;
		}
	}
	if (i >= na && j >= nb) return false;
	return i >= na;
}

static float fov_to_focal_length(int resolution, float degrees) {   // common_host.h
	return 0.5f * (float)resolution / tanf(0.5f * degrees * 3.14159265358979323846f / 180.0f);
}

// nerf_loader.cu:243-271; x_fov is in degrees, camera_angle_x in radians. false when the json names no focal length.
static bool read_focal_length(const Json& j, int res_x, int res_y, float& fx, float& fy) {
	auto one = [&](int resolution, const std::string& axis) -> float {
		if (j.contains(axis + "_fov")) return fov_to_focal_length(resolution, j[axis + "_fov"].as_float());
		if (j.contains("fl_" + axis)) return j["fl_" + axis].as_float();
		if (j.contains("camera_angle_" + axis)) return fov_to_focal_length(resolution, j["camera_angle_" + axis].as_float() * 180.0f / 3.14159265358979323846f);
		return 0.0f;
	};
	const float x_fl = one(res_x, "x"), y_fl = one(res_y, "y");
	if (x_fl != 0.0f) { fx = x_fl; fy = y_fl != 0.0f ? y_fl : x_fl; return true; }
	if (y_fl != 0.0f) { fx = fy = y_fl; return true; }
	return false;
}

struct LensDesc {
	int mode = kLensPerspective;
	float params[7] = {};
	float principal[2] = {0.5f, 0.5f};
};

// nerf_loader.cu:175-241 for the parameters the NeRF path supports
static void read_lens(const Json& j, LensDesc& lens) {
	if (j.value("is_fisheye", false) || j.contains("ftheta_p0") || j.contains("latlong") || j.contains("equirectangular")) {
		throw std::runtime_error{"ngp-b200: lens mode must be Perspective or OpenCV"};
	}
	if (j.contains("rolling_shutter")) {
		for (const Json& v : j["rolling_shutter"].elements()) if (v.as_float() != 0.0f) throw std::runtime_error{"ngp-b200: rolling shutter is outside the implemented path"};
	}
	static const struct { const char* name; int idx; } keys[] = {{"k1", 0}, {"k2", 1}, {"k3", 2}, {"k4", 3}, {"p1", 2}, {"p2", 3}};
	for (const auto& k : keys) {
		if (j.contains(k.name)) {
			lens.params[k.idx] = j[k.name].as_float();
			if (lens.params[k.idx] != 0.0f) lens.mode = kLensOpenCV;
		}
	}
	if (j.contains("cx")) lens.principal[0] = j["cx"].as_float() / j["w"].as_float();
	if (j.contains("cy")) lens.principal[1] = j["cy"].as_float() / j["h"].as_float();
}

// nerf_loader.h:101-120: flip the y / z columns, scale + offset the origin, then either cycle the axes xyz <- yzx or,
// for Mitsuba data, negate columns 0 and 2 instead. in / out: 3x4 row-major.
static void nerf_matrix_to_ngp(const float* m, float scale, const float* offset, bool from_mitsuba, float* out) {
	float r[12];
	for (int row = 0; row < 3; ++row) {
		r[row * 4 + 0] = m[row * 4 + 0];
		r[row * 4 + 1] = -m[row * 4 + 1];
		r[row * 4 + 2] = -m[row * 4 + 2];
		r[row * 4 + 3] = m[row * 4 + 3] * scale + offset[row];
	}
	if (from_mitsuba) {
		for (int row = 0; row < 3; ++row) { r[row * 4 + 0] = -r[row * 4 + 0]; r[row * 4 + 2] = -r[row * 4 + 2]; }
		for (int k = 0; k < 12; ++k) out[k] = r[k];
		return;
	}
	static const int cycle[3] = {1, 2, 0};
	for (int row = 0; row < 3; ++row) for (int c = 0; c < 4; ++c) out[row * 4 + c] = r[cycle[row] * 4 + c];
}

static const char* kImageFormats[] = {"png", "jpg", "jpeg", "bmp", "gif", "tga", "pic", "pnm", "psd", "exr"};   // nerf_loader.cu:315-317

static std::string resolve_image_path(const std::string& base, const std::string& local) {
	std::string path = path_join(base, local);
	if (path_extension(path).empty() && !file_exists(path)) {
		for (const char* ext : kImageFormats) if (file_exists(path + "." + ext)) return path + "." + ext;
	}
	return path;
}

// Testbed::load_training_data / load_nerf (src/testbed_nerf.cu:2089-2140): a json file, or a directory searched for *.json
static std::vector<std::string> json_paths(const std::string& path) {
	if (!is_directory(path)) return {path};
	std::vector<std::string> found;
	if (DIR* d = opendir(path.c_str())) {
		while (dirent* e = readdir(d)) {
			const std::string name = e->d_name;
			if (path_extension(name) == "json") found.push_back(path_join(path, name));
		}
		closedir(d);
	}
	if (found.empty()) throw std::runtime_error{"No training images were found for NeRF training!"};
	std::sort(found.begin(), found.end());
	return found;
}

struct FrameDesc {
	std::string path, alpha_path, mask_path;
	float xform_nerf[12];
	LensDesc lens;
	const Json* file_json;
	Json frame_json;
};

struct DatasetDesc {
	DatasetInfo info;
	int aabb_scale = 1;
	bool white_transparent = false, black_transparent = false, fix_premult = false;
	std::vector<FrameDesc> frames;
};

static void read_matrix_3x4(const Json& m, float* out) {
	for (int r = 0; r < 3; ++r) for (int c = 0; c < 4; ++c) out[r * 4 + c] = m[r][c].as_float();
}

// The dataset record of load_nerf without the pixels. `jsons` must outlive the result (frames point into it).
static DatasetDesc describe_dataset(const std::vector<std::string>& jsonpaths, std::vector<Json>& jsons) {
	if (jsonpaths.empty()) throw std::runtime_error{"Cannot load NeRF data from an empty set of paths."};
	jsons.clear();
	for (const std::string& p : jsonpaths) {
		const std::vector<uint8_t> text = read_file(p);
		jsons.push_back(Json::parse(std::string{text.begin(), text.end()}));
	}
	if (jsons[0].contains("camera") && jsons[0]["camera"].is_array()) throw std::runtime_error{"hdf5 is no longer supported. please use the hdf52nerf.py conversion script"};

	DatasetDesc out;
	// pass 1 (:331-396): order, cull
	std::vector<std::vector<Json>> frames_of(jsons.size());
	size_t total = 0;
	for (size_t k = 0; k < jsons.size(); ++k) {
		const Json& j = jsons[k];
		if (!j.contains("frames") || !j["frames"].is_array()) continue;
		const std::string base = path_dirname(jsonpaths[k]);
		std::vector<Json> frames = j["frames"].elements();
		std::stable_sort(frames.begin(), frames.end(), [](const Json& a, const Json& b) { return natural_less(a["file_path"].as_string(), b["file_path"].as_string()); });
		for (Json& f : frames) {
			std::string fp = f["file_path"].as_string();
			std::replace(fp.begin(), fp.end(), '\\', '/');
			f["file_path"] = fp;
		}
		if (j.contains("n_frames")) frames.resize(std::min<size_t>(frames.size(), (size_t)j["n_frames"].as_uint()));
		if (!frames.empty() && frames[0].contains("sharpness")) {
			const float thresh = j.value("sharpness_discard_threshold", 0.0f);
			std::vector<Json> kept;
			const int hood = 3, n = (int)frames.size();
			for (int i = 0; i < n; ++i) {
				const int lo = std::max(0, i - hood), hi = std::min(i + hood, n - 1);
				float mean = 0.0f;
				for (int q = lo; q < hi; ++q) mean += frames[q].value("sharpness", 1.0f);
				mean = hi > lo ? mean / (float)(hi - lo) : INFINITY;
				if (file_exists(resolve_image_path(base, frames[i]["file_path"].as_string())) && frames[i].value("sharpness", 1.0f) > thresh * mean) kept.push_back(frames[i]);
			}
			frames = kept;
		}
		total += frames.size();
		frames_of[k] = std::move(frames);
	}
	if (total == 0) throw std::runtime_error{"No training images were found for NeRF training!"};

	// pass 2 (:422-718): global settings, then the frames
	for (size_t k = 0; k < jsons.size(); ++k) {
		const Json& j = jsons[k];
		const std::string base = path_dirname(jsonpaths[k]);
		DatasetInfo& info = out.info;
		if (j.contains("normal_mts_args")) info.from_mitsuba = true;
		if (j.contains("fix_premult")) out.fix_premult = j["fix_premult"].as_bool();
		if (info.from_mitsuba) {
			info.scale = 0.66f;
			for (float& o : info.offset) o = 0.25f * 0.66f;
		}
		if (j.contains("render_aabb")) {
			info.has_render_aabb = true;
			for (int c = 0; c < 3; ++c) { info.render_aabb[c] = j["render_aabb"][0][c].as_float(); info.render_aabb[3 + c] = j["render_aabb"][1][c].as_float(); }
		}
		if (j.contains("white_transparent")) out.white_transparent = j["white_transparent"].as_bool();
		if (j.contains("black_transparent")) out.black_transparent = j["black_transparent"].as_bool();
		if (j.contains("scale")) info.scale = j["scale"].as_float();
		if (j.contains("n_extra_learnable_dims") && j["n_extra_learnable_dims"].as_int() != 0) throw std::runtime_error{"ngp-b200: per-image latents (n_extra_learnable_dims) are outside the implemented path"};
		LensDesc lens;
		read_lens(j, lens);
		if (j.contains("aabb_scale")) out.aabb_scale = j["aabb_scale"].as_int();
		if (j.contains("offset")) {
			const Json& o = j["offset"];
			for (int c = 0; c < 3; ++c) info.offset[c] = o.is_array() ? o[c].as_float() : o.as_float();
		}
		if (j.contains("aabb")) {
			float lo[3], hi[3], length = 0.000001f;
			for (int c = 0; c < 3; ++c) { lo[c] = j["aabb"][0][c].as_float(); hi[c] = j["aabb"][1][c].as_float(); length = std::max(length, fabsf(hi[c] - lo[c])); }
			info.scale = 1.0f / length;
			for (int c = 0; c < 3; ++c) info.offset[c] = ((hi[c] + lo[c]) * 0.5f) * -info.scale + 0.5f;
		}
		if (j.contains("up")) { info.up[0] = j["up"][1].as_float(); info.up[1] = j["up"][2].as_float(); info.up[2] = j["up"][0].as_float(); }   // axes permuted as the xforms
		if (!j.contains("frames") || !j["frames"].is_array()) continue;
		std::string stem = path_basename(jsonpaths[k]);
		stem = stem.substr(0, stem.find_last_of('.'));
		const std::string part_after_underscore = stem.find('_') == std::string::npos ? stem : stem.substr(stem.find_last_of('_') + 1);
		for (size_t i = 0; i < frames_of[k].size(); ++i) {
			const Json& f = frames_of[k][i];
			const Json& m = f.contains("transform_matrix_start") ? f["transform_matrix_start"] : f["transform_matrix"];
			if (f.contains("transform_matrix_end")) {
				float a[12], b[12];
				read_matrix_3x4(m, a); read_matrix_3x4(f["transform_matrix_end"], b);
				if (!std::equal(a, a + 12, b)) throw std::runtime_error{"ngp-b200: per-frame camera motion (transform_matrix_end) is outside the implemented path"};
			}
			char generated[256];
			snprintf(generated, sizeof(generated), "%s_%03d/rgba.png", part_after_underscore.c_str(), (int)i);
			const std::string rel = f["file_path"].as_string().empty() ? std::string{generated} : f["file_path"].as_string();
			FrameDesc fr;
			fr.path = resolve_image_path(base, rel);
			if (!file_exists(fr.path)) throw std::runtime_error{"Could not find image file '" + fr.path + "'."};
			const std::string ext = path_extension(fr.path);
			const std::string alpha = resolve_image_path(base, f["file_path"].as_string() + ".alpha." + ext);
			if (file_exists(alpha)) fr.alpha_path = alpha;
			std::string image_stem = path_basename(fr.path);
			image_stem = image_stem.substr(0, image_stem.find_last_of('.'));
			const std::string mask = path_join(path_dirname(fr.path), "dynamic_mask_" + image_stem + ".png");
			if (file_exists(mask)) fr.mask_path = mask;
			fr.lens = lens;
			read_lens(f, fr.lens);   // per-frame override (:703)
			read_matrix_3x4(m, fr.xform_nerf);
			fr.file_json = &jsons[k];
			fr.frame_json = f;
			if (ext == "exr") info.is_hdr = true;
			info.paths.push_back(fr.path);
			out.frames.push_back(std::move(fr));
		}
	}
	return out;
}

// ---- pixels ---------------------------------------------------------------------------------------------------
#define NGPB_NVJPEG_CHECK(x) \
	do { nvjpegStatus_t _s = (x); if (_s != NVJPEG_STATUS_SUCCESS) throw std::runtime_error{std::string{#x " failed with nvjpeg status "} + std::to_string((int)_s)}; } while (0)

// interleaved RGB (nvjpeg) -> RGBA8 texels, alpha 255
__global__ void rgb_to_rgba_kernel(uint32_t n, const uint8_t* __restrict__ rgb, uint32_t* __restrict__ rgba) {
	const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
	if (i >= n) return;
	rgba[i] = (uint32_t)rgb[3 * i] | ((uint32_t)rgb[3 * i + 1] << 8) | ((uint32_t)rgb[3 * i + 2] << 16) | 0xFF000000u;
}

NGPB_D float srgb_to_linear_host_rule(float s) { return s <= 0.04045f ? s / 12.92f : powf((s + 0.055f) / 1.055f, 2.4f); }

// The byte-image fix-ups of nerf_loader.cu:576-618 / convert_rgba32 :41-63, in the reference's order: alpha from a
// separate image (sRGB red channel -> linear), dynamic-mask pixels -> the masked-away colour 0x00FF00FF, then
// white / black pixels made transparent.
__global__ void fixup_rgba_kernel(uint32_t n, uint32_t* __restrict__ rgba, const uint32_t* __restrict__ alpha, const uint32_t* __restrict__ mask,
	bool white_transparent, bool black_transparent)
{
	const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
	if (i >= n) return;
	uint32_t v = rgba[i];
	if (alpha) {
		const float a = 255.0f * srgb_to_linear_host_rule((float)(alpha[i] & 0xFFu) * (1.0f / 255.0f));
		v = (v & 0x00FFFFFFu) | ((uint32_t)(uint8_t)a << 24);
	}
	if (mask && (mask[i] & 0x00FFFFFFu) != 0) v = 0x00FF00FFu;
	if (white_transparent && (v & 0x00FFFFFFu) == 0x00FFFFFFu) v &= 0x00FFFFFFu;
	if (black_transparent && (v & 0x00FFFFFFu) == 0u) v &= 0x00FFFFFFu;
	rgba[i] = v;
}

namespace {
struct JpegDecoder {
	nvjpegHandle_t handle = nullptr;
	nvjpegJpegState_t state = nullptr;
	~JpegDecoder() {
		if (state) nvjpegJpegStateDestroy(state);
		if (handle) nvjpegDestroy(handle);
	}
	void init() {
		if (handle) return;
		NGPB_NVJPEG_CHECK(nvjpegCreateSimple(&handle));
		NGPB_NVJPEG_CHECK(nvjpegJpegStateCreate(handle, &state));
	}
};
} // namespace

// One image file as RGBA8 texels in device memory, decoder chosen by content. JPEG: nvjpeg decodes on the GPU; PNG / BMP / TGA / PNM: host decode + upload.
static void decode_to_device(JpegDecoder& jpeg, const std::string& path, DeviceBuffer& rgb_scratch, DeviceBuffer& out, int& width, int& height, cudaStream_t stream) {
	const std::string ext = path_extension(path);
	const std::vector<uint8_t> file = read_file(path);
	// NGPB_JPEG_DECODER=host: JPEG on the host with stb_image's arithmetic (host_jpeg.cu) - the reference loader's texels bit for
	// bit, at host speed; the default decodes on the device with nvjpeg, whose IDCT and upsampling round differently (about one
	// level, tests/test_host_files_gpu.py)
	static const bool host_jpeg = getenv("NGPB_JPEG_DECODER") && std::string{getenv("NGPB_JPEG_DECODER")} == "host";
	if (is_jpeg_file(file) && !host_jpeg) {
		jpeg.init();
		int n_components = 0, widths[NVJPEG_MAX_COMPONENT], heights[NVJPEG_MAX_COMPONENT];
		nvjpegChromaSubsampling_t subsampling;
		if (nvjpegGetImageInfo(jpeg.handle, file.data(), file.size(), &n_components, &subsampling, widths, heights) != NVJPEG_STATUS_SUCCESS) {
			throw std::runtime_error{"Could not open image file: " + path};
		}
		width = widths[0]; height = heights[0];
		const size_t n = (size_t)width * height;
		rgb_scratch.reserve(n * 3, stream);
		out.alloc(n * 4, stream);
		nvjpegImage_t img{};
		img.channel[0] = rgb_scratch.as<unsigned char>();
		img.pitch[0] = (size_t)width * 3;
		NGPB_NVJPEG_CHECK(nvjpegDecode(jpeg.handle, jpeg.state, file.data(), file.size(), NVJPEG_OUTPUT_RGBI, &img, stream));
		rgb_to_rgba_kernel<<<div_round_up((uint32_t)n, 256u), 256, 0, stream>>>((uint32_t)n, rgb_scratch.as<uint8_t>(), out.as<uint32_t>());
		NGPB_CUDA_CHECK(cudaGetLastError());
		return;
	}
	if (ext != "exr") {
		const HostImage img = decode_image_rgba8(file, path);   // PNG, BMP, TGA, PNM by content (host_image.cu)
		width = img.width; height = img.height;
		out.alloc(img.rgba.size(), stream);
		NGPB_CUDA_CHECK(cudaMemcpyAsync(out.ptr, img.rgba.data(), img.rgba.size(), cudaMemcpyHostToDevice, stream));
		NGPB_CUDA_CHECK(cudaStreamSynchronize(stream));   // the host image goes away
		return;
	}
	throw std::runtime_error{"ngp-b200: '" + path + "' is an EXR file where an 8-bit image is expected"};
}

// from_rgba32 (src/common_host.cu:268-298): sRGB8 straight alpha -> linear float, colours multiplied by alpha
__global__ void rgba8_to_linear_premultiplied_kernel(uint32_t n, const uint32_t* __restrict__ rgba8, float4* __restrict__ out) {
	const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
	if (i >= n) return;
	const uint32_t v = rgba8[i];
	const float a = (float)(v >> 24) * (1.0f / 255.0f);
	out[i] = make_float4(srgb_to_linear_host_rule((float)(v & 0xFFu) * (1.0f / 255.0f)) * a, srgb_to_linear_host_rule((float)((v >> 8) & 0xFFu) * (1.0f / 255.0f)) * a,
		srgb_to_linear_host_rule((float)((v >> 16) & 0xFFu) * (1.0f / 255.0f)) * a, a);
}

// Testbed::load_image (src/testbed_image.cu:362-428): *.exr -> float RGBA as stored; 8-bit files through from_rgba32
void Testbed::load_image_file(const std::string& path) {
	const std::string ext = path_extension(path);
	if (ext == "exr") {
		const HostImageF img = decode_exr(read_file(path), path, ExrChannelRule::Image);
		set_image(img.rgba.data(), kImageFloat, img.width, img.height);
		return;
	}
	if (is_hdr_file(read_file(path))) {   // load_stbi_gpu (src/common_host.cu:268-299): Radiance files, told by content, keep their floats (stbi_loadf)
		const HostImageF img = decode_hdr(read_file(path), path);
		set_image(img.rgba.data(), kImageFloat, img.width, img.height);
		return;
	}
	JpegDecoder jpeg;
	DeviceBuffer rgb_scratch, rgba8, linear;
	int w = 0, h = 0;
	decode_to_device(jpeg, path, rgb_scratch, rgba8, w, h, m_stream);
	if (w == 0 || h == 0) throw std::runtime_error{"Image has zero pixels."};
	const uint32_t n = (uint32_t)w * (uint32_t)h;
	linear.alloc((size_t)n * sizeof(float4), m_stream);
	rgba8_to_linear_premultiplied_kernel<<<div_round_up(n, 256u), 256, 0, m_stream>>>(n, rgba8.as<uint32_t>(), linear.as<float4>());
	NGPB_CUDA_CHECK(cudaGetLastError());
	NGPB_CUDA_CHECK(cudaStreamSynchronize(m_stream));
	set_image(linear.ptr, kImageFloat, w, h);
}

void Testbed::load_training_data(const std::string& path, uint32_t max_images) {
	const std::string ext = path_extension(path);
	if (ext == "obj" || ext == "stl" || ext == "nvdb") throw std::runtime_error{"ngp-b200: SDF and volume modes are out of scope (DESIGN.md section 7)"};
	if (!file_exists(path)) throw std::runtime_error{"Data path '" + path + "' does not exist."};
	if (!is_directory(path) && ext != "json") {
		// mode_from_scene (src/testbed.cu:180-200): an image file selects the neural-image mode
		if (n_images() > 0) throw std::runtime_error{"ngp-b200: this testbed holds a NeRF scene; create a new Testbed for an image"};
		load_image_file(path);
		return;
	}
	if (m_mode == kModeImage) throw std::runtime_error{"ngp-b200: this testbed is in image mode; create a new Testbed for a NeRF scene"};

	std::vector<Json> jsons;
	DatasetDesc ds = describe_dataset(json_paths(path), jsons);
	if (max_images && ds.frames.size() > max_images) { ds.frames.resize(max_images); ds.info.paths.resize(max_images); }

	create_empty_nerf_dataset((uint32_t)ds.frames.size(), ds.aabb_scale, ds.info.is_hdr);
	JpegDecoder jpeg;
	DeviceBuffer rgb_scratch, pixels, alpha, mask;
	for (size_t i = 0; i < ds.frames.size(); ++i) {
		const FrameDesc& fr = ds.frames[i];
		int w = 0, h = 0;
		if (path_extension(fr.path) == "exr") {
			// load_exr_to_gpu (src/tinyexr_wrapper.cu): linear RGBA halves; fix_premult turns straight colours into premultiplied ones
			const HostImageF img = decode_exr(read_file(fr.path), fr.path, ExrChannelRule::Dataset);
			w = img.width; h = img.height;
			std::vector<__half> texels((size_t)w * h * 4);
			for (size_t p = 0; p < (size_t)w * h; ++p) {
				const float a = img.rgba[p * 4 + 3];
				for (int c = 0; c < 3; ++c) texels[p * 4 + c] = __float2half(ds.fix_premult ? img.rgba[p * 4 + c] * a : img.rgba[p * 4 + c]);
				texels[p * 4 + 3] = __float2half(a);
			}
			set_nerf_image((uint32_t)i, texels.data(), kImageHalf, w, h);
		} else {
		decode_to_device(jpeg, fr.path, rgb_scratch, pixels, w, h, m_stream);
		const uint32_t n = (uint32_t)w * (uint32_t)h;
		const uint32_t* alpha_px = nullptr;
		const uint32_t* mask_px = nullptr;
		if (!fr.alpha_path.empty()) {
			int aw = 0, ah = 0;
			decode_to_device(jpeg, fr.alpha_path, rgb_scratch, alpha, aw, ah, m_stream);
			if (aw != w || ah != h) throw std::runtime_error{"Alpha image " + fr.alpha_path + " has wrong resolution."};
			alpha_px = alpha.as<uint32_t>();
		}
		if (!fr.mask_path.empty()) {
			int mw = 0, mh = 0;
			decode_to_device(jpeg, fr.mask_path, rgb_scratch, mask, mw, mh, m_stream);
			if (mw != w || mh != h) throw std::runtime_error{"Dynamic mask " + fr.mask_path + " has wrong resolution."};
			mask_px = mask.as<uint32_t>();
		}
		if (alpha_px || mask_px || ds.white_transparent || ds.black_transparent) {
			fixup_rgba_kernel<<<div_round_up(n, 256u), 256, 0, m_stream>>>(n, pixels.as<uint32_t>(), alpha_px, mask_px, ds.white_transparent, ds.black_transparent);
			NGPB_CUDA_CHECK(cudaGetLastError());
		}
		NGPB_CUDA_CHECK(cudaStreamSynchronize(m_stream));
		set_nerf_image((uint32_t)i, pixels.ptr, kImageByte, w, h);
		}

		// focal length once the resolution is known (:686-690): file-level keys, then per-frame keys
		float fx = 0.0f, fy = 0.0f, pfx, pfy;
		bool have = read_focal_length(*fr.file_json, w, h, fx, fy);
		if (read_focal_length(fr.frame_json, w, h, pfx, pfy)) { fx = pfx; fy = pfy; have = true; }
		if (!have) throw std::runtime_error{"Couldn't read fov."};
		float xform[12];
		// the xforms use the FINAL scale / offset (a single transforms.json - the usual case - gives the reference's result)
		nerf_matrix_to_ngp(fr.xform_nerf, ds.info.scale, ds.info.offset, ds.info.from_mitsuba, xform);
		float lens[7] = {};
		if (fr.lens.mode == kLensOpenCV) for (int k = 0; k < 4; ++k) lens[k] = fr.lens.params[k];
		set_nerf_camera((uint32_t)i, xform, fx, fy, fr.lens.principal[0], fr.lens.principal[1], fr.lens.mode, lens);
	}
	dataset_info = ds.info;
	for (int k = 0; k < 3; ++k) view.up_dir[k] = ds.info.up[k];
}

} // namespace ngpb

// ---- host-only view of the loader's decisions (no device, no pixels): what the CPU tests compare with loader.py ----------
#include "../../include/ngp_b200.h"

extern "C" int ngpb_host_natural_less(const char* a, const char* b) {
	return a && b && ngpb::natural_less(a, b) ? 1 : 0;
}

extern "C" size_t ngpb_host_describe_dataset(const char* path, char* buf, size_t cap) {
	using namespace ngpb;
	try {
		if (!path) throw std::runtime_error{"ngpb_host_describe_dataset: null argument"};
		std::vector<Json> jsons;
		const DatasetDesc ds = describe_dataset(json_paths(path), jsons);
		auto vec = [](const float* v, int n) { Json a = Json::array(); for (int i = 0; i < n; ++i) a.push_back(Json{(double)v[i]}); return a; };
		Json j = Json::object();
		j["n_images"] = (uint32_t)ds.frames.size();
		j["aabb_scale"] = ds.aabb_scale;
		j["scale"] = (double)ds.info.scale;
		j["offset"] = vec(ds.info.offset, 3);
		j["up"] = vec(ds.info.up, 3);
		j["from_mitsuba"] = ds.info.from_mitsuba;
		j["is_hdr"] = ds.info.is_hdr;
		j["white_transparent"] = ds.white_transparent;
		j["black_transparent"] = ds.black_transparent;
		j["fix_premult"] = ds.fix_premult;
		if (ds.info.has_render_aabb) j["render_aabb"] = vec(ds.info.render_aabb, 6);
		Json frames = Json::array();
		for (const FrameDesc& fr : ds.frames) {
			Json f = Json::object();
			f["path"] = fr.path;
			f["alpha_path"] = fr.alpha_path;
			f["mask_path"] = fr.mask_path;
			float xform[12];
			nerf_matrix_to_ngp(fr.xform_nerf, ds.info.scale, ds.info.offset, ds.info.from_mitsuba, xform);
			f["xform"] = vec(xform, 12);
			f["lens_mode"] = fr.lens.mode;
			f["lens_params"] = vec(fr.lens.params, 7);
			f["principal"] = vec(fr.lens.principal, 2);
			// focal length at a nominal resolution given by the caller's json ("w", "h") when present, else 0: the real one needs the decoded image
			const int w = fr.file_json->value("w", 0), h = fr.file_json->value("h", 0);
			float fx = 0.0f, fy = 0.0f, pfx, pfy;
			bool have = w > 0 && h > 0 && read_focal_length(*fr.file_json, w, h, fx, fy);
			if (w > 0 && h > 0 && read_focal_length(fr.frame_json, w, h, pfx, pfy)) { fx = pfx; fy = pfy; have = true; }
			const float focal[2] = {have ? fx : 0.0f, have ? fy : 0.0f};
			f["focal_at_json_resolution"] = vec(focal, 2);
			frames.push_back(f);
		}
		j["frames"] = frames;
		const std::string text = j.dump();
		if (buf && cap > 0) { const size_t n = text.size() < cap - 1 ? text.size() : cap - 1; memcpy(buf, text.data(), n); buf[n] = 0; }
		return text.size() + 1;
	} catch (const std::exception& e) { ngpb::set_last_error(e.what()); return 0; }
}
