// A C++ host for libngp_b200.so with no Python in the process: what the reference's own `instant-ngp scene --network
// config.json` does headless (src/main.cu:60-200 -> Testbed::load_training_data, reload_network_from_file, train,
// save_snapshot), then a frame of training view 0 written as a binary PPM.
//
//   ngp_host <scene dir | transforms.json> <config.json> <steps> <snapshot.ingp> [frame.ppm]
//
// Build (build.py does it): g++ -std=c++17 -O2 examples/ngp_host.cpp -Iinclude -I/usr/local/cuda/include
//                                -Linstant-ngp-rendering_b200 -lngp_b200 -L/usr/local/cuda/lib64 -lcudart -o examples/ngp_host
#include <ngp_b200.h>

#include <cuda_runtime.h>

#include <chrono>
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <vector>

static void check(int rc, const char* what) {
	if (rc != NGPB_OK) {
		fprintf(stderr, "%s failed: %s\n", what, ngpb_last_error());
		exit(1);
	}
}
#define CHECK(x) check((x), #x)

int main(int argc, char** argv) {
	if (argc < 5) {
		fprintf(stderr, "usage: %s <scene> <config.json> <steps> <snapshot.ingp> [frame.ppm]\n", argv[0]);
		return 2;
	}
	const int steps = atoi(argv[3]);
	ngpb_testbed_t* t = nullptr;
	CHECK(ngpb_testbed_create(&t));
	CHECK(ngpb_testbed_load_training_data(t, argv[1], 0));
	CHECK(ngpb_testbed_reload_network_from_file(t, argv[2]));

	double n_images = 0, n_params = 0;
	CHECK(ngpb_testbed_get(t, "n_images", &n_images));
	CHECK(ngpb_testbed_get(t, "n_params", &n_params));
	printf("scene: %d views, network: %.0f parameters\n", (int)n_images, n_params);

	const auto t0 = std::chrono::steady_clock::now();
	double samples = 0;
	for (int i = 0; i < steps; ++i) {
		CHECK(ngpb_testbed_train(t, 1u << 18));
		double measured = 0;
		CHECK(ngpb_testbed_get(t, "measured_batch_size", &measured));
		samples += measured;
	}
	cudaDeviceSynchronize();
	const double seconds = std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count();
	double loss = 0;
	CHECK(ngpb_testbed_get(t, "loss", &loss));
	printf("trained %d steps in %.3f s (%.1f M samples/s), loss %.5f\n", steps, seconds, samples / seconds * 1e-6, loss);

	CHECK(ngpb_testbed_save_snapshot(t, argv[4], /*include_optimizer_state*/ 0, /*compress*/ 1));
	printf("wrote %s\n", argv[4]);

	if (argc > 5) {
		// training view 0 at a quarter of its resolution, through the view's own lens (set_camera_to_training_view, src/testbed.cu:472-487)
		ngpb_nerf_image_t view{};
		CHECK(ngpb_testbed_image_metadata(t, 0, &view));
		const uint32_t w = (uint32_t)view.res_x / 4, h = (uint32_t)view.res_y / 4;
		float camera[12];   // 3x4 rows out of the column-major mat4x3
		for (int r = 0; r < 3; ++r) for (int c = 0; c < 4; ++c) camera[r * 4 + c] = view.xform[c * 3 + r];
		CHECK(ngpb_testbed_set_render_lens(t, view.lens_mode, view.lens_params));
		CHECK(ngpb_testbed_set(t, "render_with_lens_distortion", 1.0));
		float* rgba = nullptr;
		cudaMalloc(&rgba, (size_t)w * h * 4 * sizeof(float));
		CHECK(ngpb_testbed_render(t, w, h, 1, /*linear*/ 0, camera, view.focal_x * 0.25f, view.focal_y * 0.25f, view.principal_x, view.principal_y, rgba));
		std::vector<float> host((size_t)w * h * 4);
		cudaMemcpy(host.data(), rgba, host.size() * sizeof(float), cudaMemcpyDeviceToHost);
		cudaFree(rgba);
		FILE* f = fopen(argv[5], "wb");
		if (!f) { fprintf(stderr, "cannot write %s\n", argv[5]); return 1; }
		fprintf(f, "P6\n%u %u\n255\n", w, h);
		for (size_t i = 0; i < (size_t)w * h; ++i) {
			unsigned char px[3];
			for (int c = 0; c < 3; ++c) px[c] = (unsigned char)std::lround(std::fmin(std::fmax(host[i * 4 + c], 0.0f), 1.0f) * 255.0f);
			fwrite(px, 1, 3, f);
		}
		fclose(f);
		printf("wrote %s (%ux%u)\n", argv[5], w, h);
	}
	ngpb_testbed_free(t);
	return 0;
}
