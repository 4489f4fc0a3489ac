// NetworkWithInputEncoding and NerfNetwork modules (see network.h).
#include "network.h"
#include "fused.h"

#include <cmath>

namespace ngpb {

static bool device_present() {
	int n = 0;
	if (cudaGetDeviceCount(&n) != cudaSuccess) { cudaGetLastError(); return false; }
	return n > 0;
}

static void check_batch(uint32_t n) {
	// tcnn object.h:130 CHECK_THROW(input.n() % BATCH_SIZE_GRANULARITY == 0)
	if (n % kBatchSizeGranularity != 0) {
		throw std::runtime_error{"ngp-b200: batch size " + std::to_string(n) + " must be a multiple of " + std::to_string(kBatchSizeGranularity)};
	}
}

MlpShape parse_mlp(const Json& network, uint32_t in_width, uint32_t out_width) {
	const std::string otype = network.value("otype", "MLP");
	const bool known = equals_case_insensitive(otype, "FullyFusedMLP") || equals_case_insensitive(otype, "MegakernelMLP") ||
		equals_case_insensitive(otype, "MLP") || equals_case_insensitive(otype, "CutlassMLP");
	if (!known) throw std::runtime_error{"Invalid network type: " + otype};
	const uint32_t n_neurons = network.value("n_neurons", 128u);
	if (n_neurons != kMlpWidth) {
		throw std::runtime_error{"ngp-b200 fused MLP supports n_neurons=64 (the width of every shipped NeRF / image config), but got " + std::to_string(n_neurons)};
	}
	MlpShape s;
	s.in_width = in_width;
	s.n_hidden = network.value("n_hidden_layers", 5u);
	if (s.n_hidden < 1) throw std::runtime_error{"FullyFusedMLP requires at least 1 hidden layer (3 layers in total)."};
	s.out_width = out_width;
	if (out_width > kMlpPaddedOut) throw std::runtime_error{"ngp-b200 fused MLP supports at most 16 outputs, but got " + std::to_string(out_width)};
	const std::string act = network.value("activation", "ReLU");
	if (equals_case_insensitive(act, "ReLU")) s.relu = 1;
	else if (equals_case_insensitive(act, "None")) s.relu = 0;
	else throw std::runtime_error{"ngp-b200 fused MLP: activation must be ReLU or None, but got " + act};
	const std::string oact = network.value("output_activation", "None");
	if (!equals_case_insensitive(oact, "None")) throw std::runtime_error{"ngp-b200 fused MLP: output_activation must be None, but got " + oact};
	return s;
}

// Xavier-uniform on the host with the caller's generator, matrix by matrix, element by element
// (tcnn gpu_matrix.h:284-299 via fully_fused_mlp.cu:866-891).
static void xavier_init(cudaStream_t stream, Pcg32& rng, float* params_fp32, const MlpShape& s, float scale) {
	std::vector<float> host(s.n_params());
	size_t pos = 0;
	auto fill = [&](uint32_t rows, uint32_t cols) {
		const float sc = scale * std::sqrt(6.0f / (float)(rows + cols));
		for (uint32_t i = 0; i < rows * cols; ++i) host[pos++] = rng.next_float() * 2.0f * sc - sc;
	};
	fill(kMlpWidth, s.in_width);
	for (uint32_t j = 1; j < s.n_hidden; ++j) fill(kMlpWidth, kMlpWidth);
	fill(kMlpPaddedOut, kMlpWidth);
	NGPB_CUDA_CHECK(cudaMemcpyAsync(params_fp32, host.data(), host.size() * sizeof(float), cudaMemcpyHostToDevice, stream));
	NGPB_CUDA_CHECK(cudaStreamSynchronize(stream));
}

static void push_layers(std::vector<std::pair<uint32_t, uint32_t>>& v, const MlpShape& s) {
	v.emplace_back(kMlpWidth, s.in_width);
	for (uint32_t j = 1; j < s.n_hidden; ++j) v.emplace_back(kMlpWidth, kMlpWidth);
	v.emplace_back(kMlpPaddedOut, kMlpWidth);
}

// =============================== NetworkWithInputEncoding ====================================

NetworkWithInputEncodingModule::NetworkWithInputEncodingModule(uint32_t n_input_dims, uint32_t n_output_dims, const Json& encoding, const Json& network)
: m_n_input_dims{n_input_dims}, m_encoding_json{encoding}, m_network_json{network}
{
	const std::string otype = encoding.value("otype", "OneBlob");
	uint32_t enc_out;
	if (equals_case_insensitive(otype, "Identity")) {
		m_has_grid = false;
		enc_out = n_input_dims;
		m_identity_scale = encoding.value("scale", 1.0f);
		m_identity_offset = encoding.value("offset", 0.0f);
	} else if (equals_case_insensitive(otype, "Grid") || equals_case_insensitive(otype, "HashGrid") ||
		equals_case_insensitive(otype, "TiledGrid") || equals_case_insensitive(otype, "DenseGrid")) {
		m_has_grid = true;
		m_grid = make_grid_config(n_input_dims, encoding, device_present());
		enc_out = m_grid.n_output_dims();
	} else {
		throw std::runtime_error{"Invalid encoding type: " + otype};
	}
	m_enc_width = next_multiple(enc_out, 16u); // network alignment (network_with_input_encoding.h:47-52)
	m_mlp = parse_mlp(network, m_enc_width, n_output_dims);
	m_net.nerf = 0;
	m_net.a_in = m_enc_width; m_net.a_hidden = m_mlp.n_hidden; m_net.a_relu = m_mlp.relu;
	m_net.a_param_off = 0;
	fused_check_supported(m_net);
}

std::vector<std::pair<uint32_t, uint32_t>> NetworkWithInputEncodingModule::layer_sizes() const {
	std::vector<std::pair<uint32_t, uint32_t>> v;
	push_layers(v, m_mlp);
	return v;
}

void NetworkWithInputEncodingModule::initialize_params(cudaStream_t stream, Pcg32& rng, float* params_fp32, float scale) {
	xavier_init(stream, rng, params_fp32, m_mlp, scale);
	if (m_has_grid) random_uniform(stream, rng, m_grid.n_params, params_fp32 + m_mlp.n_params(), -1e-4f * scale, 1e-4f * scale);
}

Json NetworkWithInputEncodingModule::hyperparams() const {
	Json j = Json::object();
	j["otype"] = "NetworkWithInputEncoding";
	if (m_has_grid) {
		j["encoding"] = m_grid.hyperparams();
	} else {
		// IdentityEncoding::hyperparams (tcnn encodings/identity.h:171-177)
		Json id = Json::object();
		id["otype"] = "Identity";
		id["scale"] = m_identity_scale;
		id["offset"] = m_identity_offset;
		j["encoding"] = id;
	}
	Json net = Json::object();
	net["otype"] = "FullyFusedMLP";
	net["activation"] = m_mlp.relu ? "ReLU" : "None";
	net["output_activation"] = "None";
	net["n_neurons"] = kMlpWidth;
	net["n_hidden_layers"] = m_mlp.n_hidden;
	j["network"] = net;
	return j;
}

void NetworkWithInputEncodingModule::encode_tiled(cudaStream_t stream, uint32_t n, const float* input, uint32_t in_stride, const __half* params, __half* feat) {
	if (m_has_grid) {
		grid_forward(stream, m_grid.p, n, params + m_mlp.n_params(), input, in_stride, feat, m_enc_width, 1.0f, true);
	} else {
		identity_encode_tiled(stream, n, input, in_stride, m_n_input_dims, m_identity_scale, m_identity_offset, feat, m_enc_width);
	}
}

void NetworkWithInputEncodingModule::inference(cudaStream_t stream, uint32_t n, const float* input, __half* output, const __half* params) {
	check_batch(n);
	if (n == 0) return;
	if (m_has_grid && fused_grid_supported(m_net, m_grid.p)) {
		FusedGridNetArgs f{};
		f.net = m_net; f.grid = m_grid.p; f.n = n; f.params = params; f.grid_params = params + m_mlp.n_params();
		f.input = input; f.in_stride = m_n_input_dims; f.out = output; f.out_stride = kMlpPaddedOut;
		DeviceBuffer tiles{fused_grid_gather_in_kernel() ? 0 : tiled_bytes(n, m_enc_width), stream};
		if (!fused_grid_gather_in_kernel()) {
			encode_tiled(stream, n, input, m_n_input_dims, params, tiles.as<__half>());
			f.feat_tiles = tiles.as<__half>();
		}
		fused_grid_inference(stream, f);
		return;
	}
	DeviceBuffer feat{tiled_bytes(n, m_enc_width), stream};
	encode_tiled(stream, n, input, m_n_input_dims, params, feat.as<__half>());
	FusedInferenceArgs a{};
	a.net = m_net; a.n = n; a.params = params; a.feat = feat.as<__half>();
	a.out = output; a.out_stride = kMlpPaddedOut;
	fused_inference(stream, a);
}

// Same kernels as inference() in its default (tile-fed) mode; the only difference is where the tiles live.
void NetworkWithInputEncodingModule::inference_with_scratch(cudaStream_t stream, uint32_t n, const float* input, __half* output, const __half* params,
	void* scratch, size_t scratch_bytes)
{
	check_batch(n);
	if (n == 0) return;
	if (!scratch || scratch_bytes < inference_scratch_bytes(n)) throw std::runtime_error{"ngp-b200: inference_with_scratch: scratch too small"};
	__half* tiles = (__half*)scratch;
	encode_tiled(stream, n, input, m_n_input_dims, params, tiles);
	if (m_has_grid && fused_grid_supported(m_net, m_grid.p)) {
		FusedGridNetArgs f{};
		f.net = m_net; f.grid = m_grid.p; f.n = n; f.params = params; f.grid_params = params + m_mlp.n_params();
		f.input = input; f.in_stride = m_n_input_dims; f.out = output; f.out_stride = kMlpPaddedOut;
		f.feat_tiles = tiles;
		fused_grid_inference(stream, f);
		return;
	}
	FusedInferenceArgs a{};
	a.net = m_net; a.n = n; a.params = params; a.feat = tiles;
	a.out = output; a.out_stride = kMlpPaddedOut;
	fused_inference(stream, a);
}

std::unique_ptr<Context> NetworkWithInputEncodingModule::forward(cudaStream_t stream, uint32_t n, const float* input, __half* output,
	const __half* params, bool /*prepare_input_gradients*/)
{
	// The backward kernel recomputes the tile's activations on chip, so the forward pass keeps
	// nothing (the reference stores every hidden layer, fully_fused_mlp.cu:839-852).
	if (output) inference(stream, n, input, output, params);
	else check_batch(n);
	return std::make_unique<Context>();
}

__global__ void identity_input_gradient_kernel(uint32_t total, uint32_t n_dims, const __half* __restrict__ dfeat, uint32_t stride, float scale, float* __restrict__ dL_dinput) {
	const uint32_t e = blockIdx.x * blockDim.x + threadIdx.x;
	if (e >= total) return;
	dL_dinput[e] = __half2float(dfeat[(size_t)(e / n_dims) * stride + e % n_dims]) * scale;
}

void NetworkWithInputEncodingModule::backward(cudaStream_t stream, const Context&, uint32_t n, float* dL_dinput, const __half* dL_doutput,
	__half* dL_dparams, const float* input, const __half* /*output*/, const __half* params)
{
	check_batch(n);
	if (n == 0) return;
	if (m_has_grid && !dL_dinput && fused_grid_train_enabled() && fused_grid_supported(m_net, m_grid.p)) {
		const uint32_t ctas = fused_grid_train_ctas(m_net, n);
		DeviceBuffer partials{(size_t)ctas * m_mlp.n_params() * sizeof(float), stream};
		DeviceBuffer dummy_grad{dL_dparams ? 0 : m_mlp.n_params() * sizeof(__half), stream};
		FusedGridNetArgs f{};
		f.net = m_net; f.grid = m_grid.p; f.n = n; f.params = params; f.grid_params = params + m_mlp.n_params();
		f.input = input; f.in_stride = m_n_input_dims; f.dL_dout = dL_doutput; f.dy_stride = kMlpPaddedOut;
		f.dw_partials = partials.as<float>();
		f.dL_dparams = dL_dparams ? dL_dparams : dummy_grad.as<__half>();
		__half* ggrad = dL_dparams ? dL_dparams + m_mlp.n_params() : nullptr;
		if (ggrad) NGPB_CUDA_CHECK(cudaMemsetAsync(ggrad, 0, (size_t)m_grid.n_params * sizeof(__half), stream));
		const bool tiles = !fused_grid_gather_in_kernel();
		DeviceBuffer feat{tiles ? tiled_bytes(n, m_enc_width) : 0, stream};
		DeviceBuffer dfeat{tiles && ggrad ? tiled_bytes(n, m_enc_width) : 0, stream};
		if (tiles) {
			encode_tiled(stream, n, input, m_n_input_dims, params, feat.as<__half>());
			f.feat_tiles = feat.as<__half>();
			f.dfeat_tiles = ggrad ? dfeat.as<__half>() : nullptr;
		} else {
			f.grid_grad = ggrad;
		}
		fused_grid_train(stream, f);
		if (tiles && ggrad) grid_backward(stream, m_grid.p, n, ggrad, input, m_n_input_dims, dfeat.as<__half>(), m_enc_width, 1.0f, true, m_ordered_batches);
		return;
	}
	const bool need_dfeat = (m_has_grid && dL_dparams) || dL_dinput;
	DeviceBuffer feat{tiled_bytes(n, m_enc_width), stream};
	DeviceBuffer dfeat{need_dfeat ? tiled_bytes(n, m_enc_width) : 0, stream};
	const uint32_t grid = fused_train_grid(m_net, n);
	DeviceBuffer partials{(size_t)grid * m_mlp.n_params() * sizeof(float), stream};
	DeviceBuffer dummy_grad{dL_dparams ? 0 : m_mlp.n_params() * sizeof(__half), stream};
	encode_tiled(stream, n, input, m_n_input_dims, params, feat.as<__half>());
	FusedTrainArgs a{};
	a.net = m_net; a.n = n; a.params = params; a.feat = feat.as<__half>();
	a.dL_dout = dL_doutput; a.dy_stride = kMlpPaddedOut;
	a.dfeat = need_dfeat ? dfeat.as<__half>() : nullptr;
	a.dw_partials = partials.as<float>();
	a.dL_dparams = dL_dparams ? dL_dparams : dummy_grad.as<__half>();
	fused_train(stream, a);
	if (m_has_grid && dL_dparams) {
		__half* ggrad = dL_dparams + m_mlp.n_params();
		NGPB_CUDA_CHECK(cudaMemsetAsync(ggrad, 0, (size_t)m_grid.n_params * sizeof(__half), stream));
		grid_backward(stream, m_grid.p, n, ggrad, input, m_n_input_dims, dfeat.as<__half>(), m_enc_width, 1.0f, true, m_ordered_batches);
	}
	if (dL_dinput) {
		DeviceBuffer aos{(size_t)n * m_enc_width * sizeof(__half), stream};
		tiled_to_aos(stream, n, dfeat.as<__half>(), m_enc_width, aos.as<__half>(), m_enc_width);
		if (m_has_grid) {
			grid_backward_input(stream, m_grid.p, n, params + m_mlp.n_params(), input, m_n_input_dims, aos.as<__half>(), m_enc_width,
				dL_dinput, m_n_input_dims, 1.0f);
		} else {
			// Identity encoding: dL/dx = scale * dL/dfeature over the unpadded columns (identity_backward, tcnn encodings/identity.h:66-85)
			const uint32_t total = n * m_n_input_dims;
			identity_input_gradient_kernel<<<div_round_up(total, 256u), 256, 0, stream>>>(total, m_n_input_dims, aos.as<__half>(), m_enc_width, m_identity_scale, dL_dinput);
			NGPB_CUDA_CHECK(cudaGetLastError());
		}
	}
}

// ===================================== NerfNetwork ===========================================

static uint32_t parse_sh_degree(const Json& dir_encoding, uint32_t n_dir_dims, uint32_t n_extra_dims) {
	const std::string otype = dir_encoding.value("otype", "OneBlob");
	const Json* sh = nullptr;
	if (equals_case_insensitive(otype, "SphericalHarmonics")) {
		sh = &dir_encoding;
	} else if (equals_case_insensitive(otype, "Composite")) {
		if (!dir_encoding.contains("nested") || !dir_encoding["nested"].is_array()) {
			throw std::runtime_error{"Must provide an array of nested encodings to CompositeEncoding."};
		}
		for (const Json& nested : dir_encoding["nested"].elements()) {
			const std::string t = nested.value("otype", "OneBlob");
			if (equals_case_insensitive(t, "SphericalHarmonics")) {
				if (sh) throw std::runtime_error{"ngp-b200 NerfNetwork: one SphericalHarmonics encoding expected in dir_encoding"};
				sh = &nested;
			} else if (!equals_case_insensitive(t, "Identity")) {
				throw std::runtime_error{"ngp-b200 NerfNetwork: dir_encoding may nest SphericalHarmonics and Identity, but got " + t};
			}
		}
	}
	if (!sh) throw std::runtime_error{"ngp-b200 NerfNetwork: dir_encoding must be SphericalHarmonics (optionally inside Composite), but got " + otype};
	if (n_dir_dims != 3) throw std::runtime_error{"Can only encode 3D directions in spherical harmonics."};
	if (n_extra_dims != 0) throw std::runtime_error{"ngp-b200 NerfNetwork: n_extra_dims (per-image latents) are not available"};
	const uint32_t degree = sh->value("degree", 4u);
	if (degree < 1) throw std::runtime_error{"SH degree must be at least 1"};
	return degree;
}

NerfNetworkModule::NerfNetworkModule(uint32_t n_pos_dims, uint32_t n_dir_dims, uint32_t n_extra_dims, uint32_t dir_offset,
	const Json& pos_encoding, const Json& dir_encoding, const Json& density_network, const Json& rgb_network)
: m_n_pos_dims{n_pos_dims}, m_n_dir_dims{n_dir_dims}, m_n_extra_dims{n_extra_dims}, m_dir_offset{dir_offset},
  m_pos_json{pos_encoding}, m_dir_json{dir_encoding}, m_density_json{density_network}, m_rgb_json{rgb_network}
{
	const std::string otype = pos_encoding.value("otype", "OneBlob");
	if (!(equals_case_insensitive(otype, "Grid") || equals_case_insensitive(otype, "HashGrid") ||
		equals_case_insensitive(otype, "TiledGrid") || equals_case_insensitive(otype, "DenseGrid"))) {
		throw std::runtime_error{"ngp-b200 NerfNetwork: pos encoding must be a grid encoding, but got " + otype};
	}
	m_grid = make_grid_config(n_pos_dims, pos_encoding, device_present());
	m_pos_width = next_multiple(m_grid.n_output_dims(), 16u);
	m_sh_degree = parse_sh_degree(dir_encoding, n_dir_dims, n_extra_dims);
	if (m_sh_degree > 4) throw std::runtime_error{"ngp-b200 NerfNetwork: SphericalHarmonics degree must be 1..4"};

	const uint32_t density_out = density_network.value("n_output_dims", 16u); // nerf_network.h:88-90
	m_density = parse_mlp(density_network, m_pos_width, density_out);
	const uint32_t dir_width = 16; // degree^2 padded to the rgb network's alignment
	m_rgb = parse_mlp(rgb_network, next_multiple(dir_width + kMlpPaddedOut, 16u), 3);

	m_net.nerf = 1;
	m_net.a_in = m_pos_width; m_net.a_hidden = m_density.n_hidden; m_net.a_relu = m_density.relu;
	m_net.b_in = m_rgb.in_width; m_net.b_hidden = m_rgb.n_hidden; m_net.b_relu = m_rgb.relu;
	m_net.a_param_off = 0;
	m_net.b_param_off = m_density.n_params();
	m_net.sh_degree = m_sh_degree;
	m_net.dir_offset = dir_offset;
	fused_check_supported(m_net);
}

std::vector<std::pair<uint32_t, uint32_t>> NerfNetworkModule::layer_sizes() const {
	std::vector<std::pair<uint32_t, uint32_t>> v;
	push_layers(v, m_density);
	push_layers(v, m_rgb);
	return v;
}

void NerfNetworkModule::initialize_params(cudaStream_t stream, Pcg32& rng, float* params_fp32, float scale) {
	// density MLP, rgb MLP, pos encoding, dir encoding (nerf_network.h:369-381)
	xavier_init(stream, rng, params_fp32, m_density, scale);
	xavier_init(stream, rng, params_fp32 + m_density.n_params(), m_rgb, scale);
	random_uniform(stream, rng, m_grid.n_params, params_fp32 + n_mlp_params(), -1e-4f * scale, 1e-4f * scale);
}

Json NerfNetworkModule::hyperparams() const {
	auto net_json = [](const MlpShape& s) {
		Json net = Json::object();
		net["otype"] = "FullyFusedMLP";
		net["activation"] = s.relu ? "ReLU" : "None";
		net["output_activation"] = "None";
		net["n_neurons"] = kMlpWidth;
		net["n_hidden_layers"] = s.n_hidden;
		return net;
	};
	Json j = Json::object();
	j["otype"] = "NerfNetwork";
	j["pos_encoding"] = m_grid.hyperparams();
	j["dir_encoding"] = m_dir_json;
	j["density_network"] = net_json(m_density);
	j["rgb_network"] = net_json(m_rgb);
	return j;
}

void NerfNetworkModule::inference_strided(cudaStream_t stream, uint32_t n, const float* input, uint32_t in_stride, __half* output, uint32_t out_stride, const __half* params) {
	if (n == 0) return;
	if (n % kTile != 0) throw std::runtime_error{"ngp-b200 NerfNetwork: batch size must be a multiple of 128"};
	if (fused_grid_supported(m_net, m_grid.p)) {
		FusedGridNetArgs f{};
		f.net = m_net; f.grid = m_grid.p; f.n = n; f.params = params; f.grid_params = params + pos_encoding_param_offset();
		f.input = input; f.in_stride = in_stride; f.out = output; f.out_stride = out_stride;
		if (!fused_grid_gather_in_kernel()) {
			m_scratch_feat.reserve(tiled_bytes(n, m_pos_width), stream);
			grid_forward(stream, m_grid.p, n, f.grid_params, input, in_stride, m_scratch_feat.as<__half>(), m_pos_width, 1.0f, true);
			f.feat_tiles = m_scratch_feat.as<__half>();
		}
		fused_grid_inference(stream, f);
		return;
	}
	m_scratch_feat.reserve(tiled_bytes(n, m_pos_width), stream);
	grid_forward(stream, m_grid.p, n, params + pos_encoding_param_offset(), input, in_stride, m_scratch_feat.as<__half>(), m_pos_width, 1.0f, true);
	FusedInferenceArgs a{};
	a.net = m_net; a.n = n; a.params = params; a.feat = m_scratch_feat.as<__half>();
	a.coords = input; a.coord_stride = in_stride;
	a.out = output; a.out_stride = out_stride;
	fused_inference(stream, a);
}

// Row j of the destination tiles = row src_index[j] of the source tiles (both in the tile layout of mlp.h, `chunks`
// 16-byte columns per row). Neighbouring rows of a ray stay neighbours, so both sides move in coalesced 16-byte lanes.
__global__ void __launch_bounds__(256) gather_tiles_kernel(uint32_t n, uint32_t chunks, const uint4* __restrict__ src_tiles, const uint32_t* __restrict__ src_index,
	uint32_t src_rows, uint4* __restrict__ dst_tiles)
{
	const uint32_t j = blockIdx.x * blockDim.x + threadIdx.x;
	if (j >= n) return;
	const uint32_t s = min(__ldg(src_index + j), src_rows - 1u);
	const uint4* sp = src_tiles + (size_t)(s / kTile) * chunks * kTile + (s % kTile);
	uint4* dp = dst_tiles + (size_t)(j / kTile) * chunks * kTile + (j % kTile);
	for (uint32_t ch = 0; ch < chunks; ++ch) dp[ch * kTile] = __ldg(sp + ch * kTile);
}

void NerfNetworkModule::inference_counted(cudaStream_t stream, uint32_t max_n, const uint32_t* n_dev, const float* input, uint32_t in_stride, __half* output, uint32_t out_stride, const __half* params) {
	if (max_n == 0) return;
	if (max_n % kTile != 0) throw std::runtime_error{"ngp-b200 NerfNetwork: batch size must be a multiple of 128"};
	if (fused_grid_supported(m_net, m_grid.p)) {
		FusedGridNetArgs f{};
		f.net = m_net; f.grid = m_grid.p; f.n = max_n; f.n_dev = n_dev; f.params = params; f.grid_params = params + pos_encoding_param_offset();
		f.input = input; f.in_stride = in_stride; f.out = output; f.out_stride = out_stride;
		if (!fused_grid_gather_in_kernel()) {
			m_scratch_feat.reserve(tiled_bytes(max_n, m_pos_width), stream);
			prof_mark(stream, "infer.grid_forward");
			grid_forward(stream, m_grid.p, max_n, f.grid_params, input, in_stride, m_scratch_feat.as<__half>(), m_pos_width, 1.0f, true, n_dev);
			f.feat_tiles = m_scratch_feat.as<__half>();
		}
		m_infer_tiles_rows = fused_grid_gather_in_kernel() ? 0u : max_n;
		prof_mark(stream, fused_grid_gather_in_kernel() ? "infer.fused_grid_mlp" : "infer.fused_mlp");
		fused_grid_inference(stream, f);
		return;
	}
	m_scratch_feat.reserve(tiled_bytes(max_n, m_pos_width), stream);
	prof_mark(stream, "infer.grid_forward");
	grid_forward(stream, m_grid.p, max_n, params + pos_encoding_param_offset(), input, in_stride, m_scratch_feat.as<__half>(), m_pos_width, 1.0f, true, n_dev);
	m_infer_tiles_rows = max_n;
	FusedInferenceArgs a{};
	a.net = m_net; a.n = max_n; a.n_dev = n_dev; a.params = params; a.feat = m_scratch_feat.as<__half>();
	a.coords = input; a.coord_stride = in_stride;
	a.out = output; a.out_stride = out_stride;
	prof_mark(stream, "infer.fused_mlp");
	fused_inference(stream, a);
}

void NerfNetworkModule::density(cudaStream_t stream, uint32_t n, const float* input, uint32_t in_stride, __half* output, uint32_t out_stride, const __half* params) {
	if (n == 0) return;
	if (n % kTile != 0) throw std::runtime_error{"ngp-b200 NerfNetwork: batch size must be a multiple of 128"};
	m_infer_tiles_rows = 0;   // m_scratch_feat is about to hold this call's tiles
	{
		FusedNet density_only = m_net;
		density_only.nerf = 0; // chain A (density network) only
		if (fused_grid_supported(density_only, m_grid.p)) {
			FusedGridNetArgs f{};
			f.net = density_only; f.grid = m_grid.p; f.n = n; f.params = params; f.grid_params = params + pos_encoding_param_offset();
			f.input = input; f.in_stride = in_stride; f.out = output; f.out_stride = out_stride;
			if (!fused_grid_gather_in_kernel()) {
				m_scratch_feat.reserve(tiled_bytes(n, m_pos_width), stream);
				grid_forward(stream, m_grid.p, n, f.grid_params, input, in_stride, m_scratch_feat.as<__half>(), m_pos_width, 1.0f, true);
				f.feat_tiles = m_scratch_feat.as<__half>();
			}
			fused_grid_inference(stream, f);
			return;
		}
	}
	m_scratch_feat.reserve(tiled_bytes(n, m_pos_width), stream);
	grid_forward(stream, m_grid.p, n, params + pos_encoding_param_offset(), input, in_stride, m_scratch_feat.as<__half>(), m_pos_width, 1.0f, true);
	FusedInferenceArgs a{};
	a.net = m_net;
	a.net.nerf = 0; // chain A (density network) only
	a.n = n; a.params = params; a.feat = m_scratch_feat.as<__half>();
	a.out = output; a.out_stride = out_stride;
	fused_inference(stream, a);
}

void NerfNetworkModule::backward_strided(cudaStream_t stream, uint32_t n, const float* input, uint32_t in_stride, const __half* dL_doutput, uint32_t dy_stride,
	__half* dL_dparams, const __half* params, cudaEvent_t after_mlp, cudaEvent_t before_mlp, const uint32_t* reuse_src)
{
	if (n == 0) return;
	if (n % kTile != 0) throw std::runtime_error{"ngp-b200 NerfNetwork: batch size must be a multiple of 128"};
	if (fused_grid_train_enabled() && fused_grid_supported(m_net, m_grid.p)) {
		const uint32_t ctas = fused_grid_train_ctas(m_net, n);
		m_scratch_dw.reserve((size_t)ctas * n_mlp_params() * sizeof(float), stream);
		FusedGridNetArgs f{};
		f.net = m_net; f.grid = m_grid.p; f.n = n; f.params = params; f.grid_params = params + pos_encoding_param_offset();
		f.input = input; f.in_stride = in_stride; f.dL_dout = dL_doutput; f.dy_stride = dy_stride;
		__half* ggrad = dL_dparams + pos_encoding_param_offset();
		f.dw_partials = m_scratch_dw.as<float>();
		f.dL_dparams = dL_dparams;
		const bool tiles = !fused_grid_gather_in_kernel();
		if (tiles) {
			m_scratch_feat.reserve(tiled_bytes(n, m_pos_width), stream);
			m_scratch_dfeat.reserve(tiled_bytes(n, m_pos_width), stream);
			m_infer_tiles_rows = 0;
			prof_mark(stream, "train.grid_forward");
			grid_forward(stream, m_grid.p, n, f.grid_params, input, in_stride, m_scratch_feat.as<__half>(), m_pos_width, 1.0f, true);
			f.feat_tiles = m_scratch_feat.as<__half>();
			f.dfeat_tiles = m_scratch_dfeat.as<__half>();
		} else {
			f.grid_grad = ggrad;
			prof_mark(stream, "train.grid_grad_memset");
			NGPB_CUDA_CHECK(cudaMemsetAsync(ggrad, 0, (size_t)m_grid.n_params * sizeof(__half), stream));
		}
		prof_mark(stream, tiles ? "train.fused_mlp_fwd_bwd" : "train.fused_grid_mlp_fwd_bwd_scatter");
		if (before_mlp) NGPB_CUDA_CHECK(cudaEventRecord(before_mlp, stream));
		fused_grid_train(stream, f);
		if (after_mlp) NGPB_CUDA_CHECK(cudaEventRecord(after_mlp, stream));
		if (tiles) {
			prof_mark(stream, "train.grid_backward");
			NGPB_CUDA_CHECK(cudaMemsetAsync(ggrad, 0, (size_t)m_grid.n_params * sizeof(__half), stream));
			grid_backward(stream, m_grid.p, n, ggrad, input, in_stride, m_scratch_dfeat.as<__half>(), m_pos_width, 1.0f, true, true);
		}
		return;
	}
	const uint32_t grid = fused_train_grid(m_net, n);
	m_scratch_dfeat.reserve(tiled_bytes(n, m_pos_width), stream);
	m_scratch_dw.reserve((size_t)grid * n_mlp_params() * sizeof(float), stream);
	static const bool reuse_off = getenv("NGPB_REUSE_TILES") && atoi(getenv("NGPB_REUSE_TILES")) == 0;
	const __half* feat;
	if (reuse_src && m_infer_tiles_rows > 0 && !reuse_off) {
		// the compacted batch's rows are copies of marched samples whose features the inference pass left behind: 64 bytes
		// copied per sample instead of 64 table entries gathered (0.039 -> 0.01 ms at 2^18 samples)
		m_scratch_feat_train.reserve(tiled_bytes(n, m_pos_width), stream);
		prof_mark(stream, "train.reuse_feature_tiles");
		gather_tiles_kernel<<<div_round_up(n, 256u), 256, 0, stream>>>(n, m_pos_width / 8, reinterpret_cast<const uint4*>(m_scratch_feat.ptr), reuse_src, m_infer_tiles_rows,
			reinterpret_cast<uint4*>(m_scratch_feat_train.ptr));
		NGPB_CUDA_CHECK(cudaGetLastError());
		feat = m_scratch_feat_train.as<__half>();
	} else {
		m_scratch_feat.reserve(tiled_bytes(n, m_pos_width), stream);
		m_infer_tiles_rows = 0;
		prof_mark(stream, "train.grid_forward");
		grid_forward(stream, m_grid.p, n, params + pos_encoding_param_offset(), input, in_stride, m_scratch_feat.as<__half>(), m_pos_width, 1.0f, true);
		feat = m_scratch_feat.as<__half>();
	}
	FusedTrainArgs a{};
	a.net = m_net; a.n = n; a.params = params; a.feat = feat;
	a.coords = input; a.coord_stride = in_stride;
	a.dL_dout = dL_doutput; a.dy_stride = dy_stride;
	a.dfeat = m_scratch_dfeat.as<__half>();
	a.dw_partials = m_scratch_dw.as<float>();
	a.dL_dparams = dL_dparams;
	prof_mark(stream, "train.fused_mlp_fwd_bwd");
	if (before_mlp) NGPB_CUDA_CHECK(cudaEventRecord(before_mlp, stream));
	fused_train(stream, a);
	if (after_mlp) NGPB_CUDA_CHECK(cudaEventRecord(after_mlp, stream));
	prof_mark(stream, "train.grid_backward");
	__half* ggrad = dL_dparams + pos_encoding_param_offset();
	NGPB_CUDA_CHECK(cudaMemsetAsync(ggrad, 0, (size_t)m_grid.n_params * sizeof(__half), stream));
	grid_backward(stream, m_grid.p, n, ggrad, input, in_stride, m_scratch_dfeat.as<__half>(), m_pos_width, 1.0f, true, true);
}

void NerfNetworkModule::inference(cudaStream_t stream, uint32_t n, const float* input, __half* output, const __half* params) {
	check_batch(n);
	inference_strided(stream, n, input, n_input_dims(), output, kMlpPaddedOut, params);
}

std::unique_ptr<Context> NerfNetworkModule::forward(cudaStream_t stream, uint32_t n, const float* input, __half* output,
	const __half* params, bool /*prepare_input_gradients*/)
{
	check_batch(n);
	if (output) inference_strided(stream, n, input, n_input_dims(), output, kMlpPaddedOut, params);
	return std::make_unique<Context>();
}

void NerfNetworkModule::backward(cudaStream_t stream, const Context&, uint32_t n, float* dL_dinput, const __half* dL_doutput,
	__half* dL_dparams, const float* input, const __half* /*output*/, const __half* params)
{
	check_batch(n);
	if (dL_dinput) throw std::runtime_error{"ngp-b200 NerfNetwork: input gradients (camera / latent optimisation) are not available"};
	if (!dL_dparams) return;
	backward_strided(stream, n, input, n_input_dims(), dL_doutput, kMlpPaddedOut, dL_dparams, params);
}

// ======================================= factories ===========================================

Module* create_network_with_input_encoding_module(uint32_t n_input_dims, uint32_t n_output_dims, const Json& encoding, const Json& network) {
	return new NetworkWithInputEncodingModule{n_input_dims, n_output_dims, encoding, network};
}

Module* create_nerf_network_module(uint32_t n_pos_dims, uint32_t n_dir_dims, uint32_t n_extra_dims, uint32_t dir_offset,
	const Json& pos_encoding, const Json& dir_encoding, const Json& density_network, const Json& rgb_network)
{
	return new NerfNetworkModule{n_pos_dims, n_dir_dims, n_extra_dims, dir_offset, pos_encoding, dir_encoding, density_network, rgb_network};
}

} // namespace ngpb
