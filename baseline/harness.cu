// GPU-oracle / GPU-baseline harness: drives the UNMODIFIED reference kernels (tiny-cuda-nn and
// instant-ngp's NerfNetwork, compiled by baseline/build_ref.sh from /root/reference) on raw binary
// inputs and writes their raw outputs, plus CUDA-event timings.  Test infrastructure and baseline only:
// nothing in the product links or loads it.  It includes the reference's headers from where they lie
// and calls its public C++ API (tcnn::cpp::Module, cpp_api.h:86-119; tcnn::create_optimizer,
// optimizer.h:66; ngp::NerfNetwork, nerf_network.h:81); nothing is copied.
//
//   ref_harness grid    <dir>   cfg.json{n_dims, encoding}           x.bin params.bin dy.bin -> y.bin dparams.bin
//   ref_harness mlp     <dir>   cfg.json{n_in, n_out, network}        x.bin params.bin dy.bin -> y.bin dx.bin dparams.bin
//   ref_harness nerfnet <dir>   cfg.json{pos/dir encoding, networks}  x.bin params.bin dy.bin -> y.bin dparams.bin density.bin
//   ref_harness adam    <dir>   cfg.json{optimizer, layers, n, steps} w.bin g.bin(steps x n)  -> w32.bin w16.bin ema.bin
// Every mode also writes timing.json {"fwd_ms": [...], "bwd_ms": [...]} when cfg.json has "time_iters" > 0
// (each iteration timed on its own with CUDA events, a 256 MB write between iterations to flush L2:
// method of tcnn benchmarks/image/bench_ours.cu:203-336 with the per-iteration flush our own bench uses).
#include <tiny-cuda-nn/common.h>
#include <tiny-cuda-nn/cpp_api.h>
#include <tiny-cuda-nn/gpu_memory.h>
#include <tiny-cuda-nn/optimizer.h>
#include <tiny-cuda-nn/encoding.h>
#include <tiny-cuda-nn/network.h>

#include <neural-graphics-primitives/common.h>
#include <neural-graphics-primitives/nerf_network.h>

#include <algorithm>
#include <cstdio>
#include <fstream>
#include <iostream>
#include <string>
#include <vector>

using json = nlohmann::json;
using namespace tcnn;

static std::vector<uint8_t> read_file(const std::string& path) {
	std::ifstream f(path, std::ios::binary | std::ios::ate);
	if (!f) throw std::runtime_error("cannot open " + path);
	std::vector<uint8_t> b((size_t)f.tellg());
	f.seekg(0);
	f.read((char*)b.data(), b.size());
	return b;
}
static void write_file(const std::string& path, const void* p, size_t n) {
	std::ofstream f(path, std::ios::binary);
	f.write((const char*)p, n);
}
template <typename T> static GPUMemory<T> upload(const std::string& path) {
	auto b = read_file(path);
	GPUMemory<T> m(b.size() / sizeof(T));
	CUDA_CHECK_THROW(cudaMemcpy(m.data(), b.data(), b.size(), cudaMemcpyHostToDevice));
	return m;
}
template <typename T> static void download(const std::string& path, const GPUMemory<T>& m, size_t n = 0) {
	if (n == 0) n = m.size();
	std::vector<T> h(n);
	CUDA_CHECK_THROW(cudaMemcpy(h.data(), m.data(), n * sizeof(T), cudaMemcpyDeviceToHost));
	write_file(path, h.data(), n * sizeof(T));
}

struct Timer {
	cudaEvent_t a, b;
	GPUMemory<uint8_t> flush{256u << 20};
	Timer() { cudaEventCreate(&a); cudaEventCreate(&b); }
	template <typename F> float run(cudaStream_t s, F&& f) {
		CUDA_CHECK_THROW(cudaMemsetAsync(flush.data(), 1, flush.size(), s));
		CUDA_CHECK_THROW(cudaEventRecord(a, s));
		f();
		CUDA_CHECK_THROW(cudaEventRecord(b, s));
		CUDA_CHECK_THROW(cudaEventSynchronize(b));
		float ms = 0;
		cudaEventElapsedTime(&ms, a, b);
		return ms;
	}
};

// forward + backward of a cpp::Module on the files of `dir`
static void run_module(cpp::Module* mod, const std::string& dir, const json& cfg, bool want_dx) {
	cudaStream_t stream;
	CUDA_CHECK_THROW(cudaStreamCreate(&stream));
	auto x = upload<float>(dir + "/x.bin");
	const uint32_t n = (uint32_t)(x.size() / mod->n_input_dims());
	const uint32_t n_out = mod->n_output_dims();
	GPUMemory<__half> params;
	if (cfg.value("init_seed", -1) >= 0) {
		// the reference's own initialiser (pcg32-driven; cpp_api.cu:134-137), dumped so that ours can be compared
		GPUMemory<float> p32(mod->n_params());
		mod->initialize_params((size_t)cfg["init_seed"].get<int>(), p32.data(), 1.0f);
		download(dir + "/init_params32.bin", p32);
		std::vector<float> h(p32.size());
		p32.copy_to_host(h);
		std::vector<__half> hh(h.size());
		for (size_t i = 0; i < h.size(); ++i) hh[i] = (__half)h[i];
		params.resize(hh.size());
		params.copy_from_host(hh);
	} else {
		params = upload<__half>(dir + "/params.bin");
	}
	if (params.size() != mod->n_params()) throw std::runtime_error("params.bin holds " + std::to_string(params.size()) + " values, the module has " + std::to_string(mod->n_params()));
	auto dy = upload<__half>(dir + "/dy.bin");
	GPUMemory<__half> y((size_t)n * n_out), dparams(params.size());
	GPUMemory<float> dx(want_dx ? x.size() : 0);
	y.memset(0);
	dparams.memset(0);
	auto ctx = mod->forward(stream, n, x.data(), y.data(), params.data(), want_dx);
	mod->backward(stream, ctx, n, want_dx ? dx.data() : nullptr, dy.data(), dparams.data(), x.data(), y.data(), params.data());
	CUDA_CHECK_THROW(cudaStreamSynchronize(stream));
	download(dir + "/y.bin", y);
	download(dir + "/dparams.bin", dparams);
	if (want_dx) download(dir + "/dx.bin", dx);

	const int iters = cfg.value("time_iters", 0);
	if (iters > 0) {
		Timer t;
		std::vector<float> fwd, bwd, inf;
		for (int i = 0; i < iters + 3; ++i) {
			cpp::Context c;
			const float f = t.run(stream, [&] { c = mod->forward(stream, n, x.data(), y.data(), params.data(), want_dx); });
			const float b = t.run(stream, [&] { mod->backward(stream, c, n, want_dx ? dx.data() : nullptr, dy.data(), dparams.data(), x.data(), y.data(), params.data()); });
			const float q = t.run(stream, [&] { mod->inference(stream, n, x.data(), y.data(), params.data()); });
			if (i >= 3) { fwd.push_back(f); bwd.push_back(b); inf.push_back(q); }
		}
		json out = {{"n", n}, {"fwd_ms", fwd}, {"bwd_ms", bwd}, {"inference_ms", inf}};
		std::ofstream(dir + "/timing.json") << out.dump();
	}
}

static void run_nerfnet(const std::string& dir, const json& cfg) {
	using T = network_precision_t;
	cudaStream_t stream;
	CUDA_CHECK_THROW(cudaStreamCreate(&stream));
	// the call Testbed::reset_network makes (src/testbed.cu:3848-3858): 3 position dims, 3 direction dims at offset 4
	const uint32_t n_in = 7, dir_offset = 4;
	auto net = std::make_shared<ngp::NerfNetwork<T>>(3, 3, 0, dir_offset, cfg["encoding"], cfg["dir_encoding"], cfg["network"], cfg["rgb_network"]);
	auto xin = upload<float>(dir + "/x.bin");   // [n][7] NerfCoordinate rows
	const uint32_t n = (uint32_t)(xin.size() / n_in);
	auto params = upload<T>(dir + "/params.bin");
	if (params.size() != net->n_params()) throw std::runtime_error("params.bin holds " + std::to_string(params.size()) + " values, NerfNetwork has " + std::to_string(net->n_params()));
	auto dy = upload<T>(dir + "/dy.bin");       // [n][16]
	GPUMemory<T> dparams(params.size());
	dparams.memset(0);
	net->set_params(params.data(), params.data(), dparams.data());
	const uint32_t n_out = net->padded_output_width();
	GPUMemory<T> y((size_t)n * n_out), dens((size_t)n * 16);
	GPUMatrix<float> input(xin.data(), n_in, n);
	GPUMatrix<T> output(y.data(), n_out, n), dL(dy.data(), n_out, n);
	net->inference_mixed_precision(stream, input, output);
	CUDA_CHECK_THROW(cudaStreamSynchronize(stream));
	download(dir + "/y_inference.bin", y);
	{
		GPUMatrix<T> d(dens.data(), 16, n);
		net->density(stream, input.slice_rows(0, 3), d);  // what update_density_grid_nerf calls (src/testbed_nerf.cu:2325)
		CUDA_CHECK_THROW(cudaStreamSynchronize(stream));
		download(dir + "/density.bin", dens);
	}
	auto ctx = net->forward(stream, input, &output, false, false);
	net->backward(stream, *ctx, input, output, dL, nullptr, false, GradientMode::Overwrite);
	CUDA_CHECK_THROW(cudaStreamSynchronize(stream));
	download(dir + "/y.bin", y);
	download(dir + "/dparams.bin", dparams);
	json info = {{"n_params", net->n_params()}, {"padded_output_width", n_out}};
	const int iters = cfg.value("time_iters", 0);
	if (iters > 0) {
		Timer t;
		std::vector<float> fb, inf;
		for (int i = 0; i < iters + 3; ++i) {
			const float a = t.run(stream, [&] {
				auto c = net->forward(stream, input, &output, false, false);
				net->backward(stream, *c, input, output, dL, nullptr, false, GradientMode::Overwrite);
			});
			const float q = t.run(stream, [&] { net->inference_mixed_precision(stream, input, output); });
			if (i >= 3) { fb.push_back(a); inf.push_back(q); }
		}
		info["n"] = n; info["fwd_bwd_ms"] = fb; info["inference_ms"] = inf;
	}
	std::ofstream(dir + "/timing.json") << info.dump();
}

static void run_adam(const std::string& dir, const json& cfg) {
	using T = network_precision_t;
	cudaStream_t stream;
	CUDA_CHECK_THROW(cudaStreamCreate(&stream));
	std::shared_ptr<Optimizer<T>> opt{create_optimizer<T>(cfg["optimizer"])};
	const uint32_t n = cfg["n"], steps = cfg["steps"];
	std::vector<std::pair<uint32_t, uint32_t>> layers;
	for (auto& l : cfg["layers"]) layers.emplace_back(l[0].get<uint32_t>(), l[1].get<uint32_t>());
	opt->allocate(n, layers);
	auto w32 = upload<float>(dir + "/w.bin");
	auto g = upload<T>(dir + "/g.bin");
	GPUMemory<T> w16(n);
	{
		std::vector<float> h(n);
		w32.copy_to_host(h);
		std::vector<T> hh(n);
		for (uint32_t i = 0; i < n; ++i) hh[i] = (T)h[i];
		w16.copy_from_host(hh);
	}
	const float loss_scale = cfg.value("loss_scale", 128.0f);
	for (uint32_t s = 0; s < steps; ++s) opt->step(stream, loss_scale, w32.data(), w16.data(), g.data() + (size_t)s * n);
	CUDA_CHECK_THROW(cudaStreamSynchronize(stream));
	download(dir + "/w32.bin", w32);
	download(dir + "/w16.bin", w16);
	if (opt->custom_weights()) {
		std::vector<T> h(n);
		CUDA_CHECK_THROW(cudaMemcpy(h.data(), opt->custom_weights(), n * sizeof(T), cudaMemcpyDeviceToHost));
		write_file(dir + "/ema.bin", h.data(), n * sizeof(T));
	}
	json info = {{"learning_rate", opt->learning_rate()}, {"step", opt->step()}};
	const int iters = cfg.value("time_iters", 0);
	if (iters > 0) {
		Timer t;
		std::vector<float> ms;
		for (int i = 0; i < iters; ++i) ms.push_back(t.run(stream, [&] { opt->step(stream, loss_scale, w32.data(), w16.data(), g.data()); }));
		info["step_ms"] = ms;
	}
	std::ofstream(dir + "/timing.json") << info.dump();
}

int main(int argc, char** argv) {
	if (argc < 3) {
		std::fprintf(stderr, "usage: ref_harness grid|mlp|nerfnet|adam <dir>\n");
		return 2;
	}
	const std::string mode = argv[1], dir = argv[2];
	try {
		json cfg = json::parse(std::ifstream(dir + "/cfg.json"));
		if (mode == "grid") {
			std::unique_ptr<cpp::Module> m{cpp::create_encoding(cfg["n_dims"], cfg["encoding"], cpp::Precision::Fp16)};
			run_module(m.get(), dir, cfg, false);
		} else if (mode == "mlp") {
			std::unique_ptr<cpp::Module> m{cpp::create_network(cfg["n_in"], cfg["n_out"], cfg["network"])};
			run_module(m.get(), dir, cfg, true);
		} else if (mode == "net") {
			std::unique_ptr<cpp::Module> m{cpp::create_network_with_input_encoding(cfg["n_dims"], cfg["n_out"], cfg["encoding"], cfg["network"])};
			run_module(m.get(), dir, cfg, false);
		} else if (mode == "nerfnet") {
			run_nerfnet(dir, cfg);
		} else if (mode == "adam") {
			run_adam(dir, cfg);
		} else {
			throw std::runtime_error("unknown mode " + mode);
		}
	} catch (const std::exception& e) {
		std::fprintf(stderr, "ref_harness: %s\n", e.what());
		return 1;
	}
	return 0;
}
