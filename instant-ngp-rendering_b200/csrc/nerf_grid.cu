// Occupancy ("density") grid maintenance: src/testbed_nerf.cu:74-331 and the host sequences at
// :2271-2379. Cells are stored in Morton order per cascade; the bitfield packs 8 consecutive
// Morton cells per byte, cascade k at byte offset k * 128^3 / 8.
#include "nerf.h"

namespace ngpb {

// pos_to_uv for perspective / OpenCV cameras (common_device.cuh:497-540): project a world point
// into the image of camera `m`; with an OpenCV lens the distortion delta is applied forward.
NGPB_D void pos_to_uv(V3 pos, const ImageMeta& m, float& u, float& v) {
	const V3 d = pos - m.xform.c[3];
	// inverse of the (orthonormal up to scale) rotation: the reference uses inverse(mat3(camera))
	const V3 c0 = m.xform.c[0], c1 = m.xform.c[1], c2 = m.xform.c[2];
	const float det = c0.x * (c1.y * c2.z - c2.y * c1.z) - c1.x * (c0.y * c2.z - c2.y * c0.z) + c2.x * (c0.y * c1.z - c1.y * c0.z);
	const float inv = 1.0f / det;
	V3 dir;
	dir.x = inv * (d.x * (c1.y * c2.z - c2.y * c1.z) - c1.x * (d.y * c2.z - c2.y * d.z) + c2.x * (d.y * c1.z - c1.y * d.z));
	dir.y = inv * (c0.x * (d.y * c2.z - c2.y * d.z) - d.x * (c0.y * c2.z - c2.y * c0.z) + c2.x * (c0.y * d.z - d.y * c0.z));
	dir.z = inv * (c0.x * (c1.y * d.z - d.y * c1.z) - c1.x * (c0.y * d.z - d.y * c0.z) + d.x * (c0.y * c1.z - c1.y * c0.z));
	dir = dir / dir.z;
	float du = 0.0f, dv = 0.0f;
	if (m.lens_mode == kLensOpenCV) opencv_lens_distortion_delta(m.lens_params, dir.x, dir.y, &du, &dv);
	dir.x += du;
	dir.y += dv;
	u = dir.x * m.focal_x / (float)m.res_x + m.principal_x;
	v = dir.y * m.focal_y / (float)m.res_y + m.principal_y;
}

__global__ void mark_untrained_density_grid_kernel(uint32_t n_elements, float* __restrict__ grid_out, uint32_t n_training_images,
	const ImageMeta* __restrict__ metadata, bool clear_visible_voxels)
{
	const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
	if (i >= n_elements) return;
	const uint32_t level = i / kNerfGridCells, pos_idx = i % kNerfGridCells;
	const uint32_t x = morton3d_invert(pos_idx >> 0), y = morton3d_invert(pos_idx >> 1), z = morton3d_invert(pos_idx >> 2);
	const float voxel_size = scalbnf(1.0f / kNerfGridSize, (int)level);
	const V3 pos = (v3((float)x, (float)y, (float)z) / (float)kNerfGridSize - 0.5f) * scalbnf(1.0f, (int)level) + 0.5f;

	const uint32_t min_count = 1;
	uint32_t count = 0;
	for (uint32_t j = 0; j < n_training_images && count < min_count; ++j) {
		const ImageMeta& m = metadata[j];
		if (m.lens_mode == kLensFTheta || m.lens_mode == kLensLatLong || m.lens_mode == kLensEquirectangular) { ++count; continue; }
		for (uint32_t k = 0; k < 8; ++k) {
			const V3 corner = pos + v3((k & 1) ? voxel_size : 0.0f, (k & 2) ? voxel_size : 0.0f, (k & 4) ? voxel_size : 0.0f);
			const V3 dir = normalize(corner - m.xform.c[3]);
			if (dot(dir, m.xform.c[2]) < 1e-4f) continue;
			float u, v;
			pos_to_uv(corner, m, u, v);
			// pos_to_uv is not injective under lens distortion: re-generate the ray and compare.
			const Ray3 ray = uv_to_ray(u, v, m.res_x, m.res_y, m.focal_x, m.focal_y, m.xform, m.principal_x, m.principal_y, m.lens_mode, m.lens_params, 0.0f);
			if (distance(normalize(ray.d), dir) < 1e-3f && u > 0.0f && v > 0.0f && u < 1.0f && v < 1.0f) { ++count; break; }
		}
	}
	if (clear_visible_voxels || (grid_out[i] < 0) != (count < min_count)) grid_out[i] = (count >= min_count) ? 0.f : -1.f;
}

__global__ void generate_grid_samples_kernel(uint32_t n_elements, Pcg32 rng, uint32_t step, Aabb aabb, const float* __restrict__ grid_in,
	float* __restrict__ out, uint32_t* __restrict__ indices, uint32_t n_cascades, float thresh)
{
	const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
	if (i >= n_elements) return;
	// 1 random number to select the level, 3 to select the position
	rng.advance((int64_t)i * 4);
	const uint32_t level = (uint32_t)(rng.next_float() * n_cascades) % n_cascades;
	uint32_t idx = 0;
	for (uint32_t j = 0; j < 10; ++j) {
		idx = ((i + step * n_elements) * 56924617u + j * 19349663u + 96925573u) % kNerfGridCells;
		idx += level * kNerfGridCells;
		if (grid_in[idx] > thresh) break;
	}
	const uint32_t pos_idx = idx % kNerfGridCells;
	const uint32_t x = morton3d_invert(pos_idx >> 0), y = morton3d_invert(pos_idx >> 1), z = morton3d_invert(pos_idx >> 2);
	const float rx = rng.next_float(), ry = rng.next_float(), rz = rng.next_float();
	const V3 pos = ((v3((float)x, (float)y, (float)z) + v3(rx, ry, rz)) / (float)kNerfGridSize - 0.5f) * scalbnf(1.0f, (int)level) + 0.5f;
	const V3 wp = warp_position(pos, aabb);
	reinterpret_cast<float4*>(out)[i] = make_float4(wp.x, wp.y, wp.z, warp_dt(kMinConeStepsize));
	indices[i] = idx;
}

__global__ void splat_grid_samples_kernel(uint32_t n_elements, const uint32_t* __restrict__ indices, const __half* __restrict__ density_raw,
	float* __restrict__ grid_out, int density_activation)
{
	const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
	if (i >= n_elements) return;
	const float mlp = network_to_density(__half2float(density_raw[i]), density_activation);
	const float optical_thickness = mlp * kMinConeStepsize;
	// positive floats order like their bit patterns
	atomicMax(reinterpret_cast<uint32_t*>(grid_out) + indices[i], __float_as_uint(optical_thickness));
}

__global__ void ema_grid_samples_kernel(uint32_t n_elements, float decay, float* __restrict__ grid_out, const float* __restrict__ grid_in) {
	const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
	if (i >= n_elements) return;
	const float importance = grid_in[i];
	const float prev_val = grid_out[i];
	grid_out[i] = (prev_val < 0.f) ? prev_val : fmaxf(prev_val * decay, importance);
}

// sum over cascade 0 of max(v, 0) / n (the reference's reduce_sum with a transform, :2369-2370)
__global__ void mean_density_kernel(uint32_t n, const float* __restrict__ grid, float* __restrict__ mean_out) {
	__shared__ float warp_part[8];
	float s = 0.0f;
	for (uint32_t i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) s += fmaxf(grid[i], 0.f) / (float)n;
#pragma unroll
	for (uint32_t o = 16; o > 0; o >>= 1) s += __shfl_xor_sync(0xFFFFFFFFu, s, o);
	if (threadIdx.x % 32 == 0) warp_part[threadIdx.x / 32] = s;
	__syncthreads();
	if (threadIdx.x == 0) {
		float t = 0.0f;
		for (uint32_t w = 0; w < blockDim.x / 32; ++w) t += warp_part[w];
		atomicAdd(mean_out, t);
	}
}

__global__ void grid_to_bitfield_kernel(uint32_t n_elements, uint32_t n_nonzero_elements, const float* __restrict__ grid, uint8_t* __restrict__ bitfield,
	const float* __restrict__ mean_density_ptr)
{
	const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
	if (i >= n_elements) return;
	if (i >= n_nonzero_elements) { bitfield[i] = 0; return; }
	const float thresh = fminf(kNerfMinOpticalThickness, *mean_density_ptr);
	const float4 a = reinterpret_cast<const float4*>(grid)[i * 2], b = reinterpret_cast<const float4*>(grid)[i * 2 + 1];
	uint8_t bits = 0;
	bits |= a.x > thresh ? 1 : 0; bits |= a.y > thresh ? 2 : 0; bits |= a.z > thresh ? 4 : 0; bits |= a.w > thresh ? 8 : 0;
	bits |= b.x > thresh ? 16 : 0; bits |= b.y > thresh ? 32 : 0; bits |= b.z > thresh ? 64 : 0; bits |= b.w > thresh ? 128 : 0;
	bitfield[i] = bits;
}

__global__ void bitfield_max_pool_kernel(uint32_t n_elements, const uint8_t* __restrict__ prev_level, uint8_t* __restrict__ next_level) {
	const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
	if (i >= n_elements) return;
	const uint2 p = reinterpret_cast<const uint2*>(prev_level)[i];
	uint8_t bits = 0;
#pragma unroll
	for (uint32_t j = 0; j < 8; ++j) {
		const uint32_t byte = ((j < 4 ? p.x : p.y) >> (8 * (j & 3))) & 0xFFu;
		bits |= byte > 0 ? (uint8_t)(1u << j) : 0;
	}
	const uint32_t x = morton3d_invert(i >> 0) + kNerfGridSize / 8;
	const uint32_t y = morton3d_invert(i >> 1) + kNerfGridSize / 8;
	const uint32_t z = morton3d_invert(i >> 2) + kNerfGridSize / 8;
	next_level[morton3d(x, y, z)] |= bits;
}

// ------------------------------------------------------------------------------------------------
void nerf_mark_untrained_density_grid(cudaStream_t stream, uint32_t n_elements, float* grid, uint32_t n_images, const ImageMeta* metadata, bool clear_visible_voxels) {
	if (!n_elements) return;
	mark_untrained_density_grid_kernel<<<div_round_up(n_elements, 128), 128, 0, stream>>>(n_elements, grid, n_images, metadata, clear_visible_voxels);
	NGPB_CUDA_CHECK(cudaGetLastError());
}
void nerf_generate_grid_samples(cudaStream_t stream, uint32_t n, Pcg32 rng, uint32_t step, const Aabb& aabb, const float* grid_in,
	float* positions, uint32_t* indices, uint32_t n_cascades, float thresh)
{
	if (!n) return;
	generate_grid_samples_kernel<<<div_round_up(n, 128), 128, 0, stream>>>(n, rng, step, aabb, grid_in, positions, indices, n_cascades, thresh);
	NGPB_CUDA_CHECK(cudaGetLastError());
}
void nerf_splat_grid_samples(cudaStream_t stream, uint32_t n, const uint32_t* indices, const __half* density_raw, float* grid_tmp, int density_activation) {
	if (!n) return;
	splat_grid_samples_kernel<<<div_round_up(n, 256), 256, 0, stream>>>(n, indices, density_raw, grid_tmp, density_activation);
	NGPB_CUDA_CHECK(cudaGetLastError());
}
void nerf_ema_grid_samples(cudaStream_t stream, uint32_t n_elements, float decay, float* grid, const float* grid_tmp) {
	if (!n_elements) return;
	ema_grid_samples_kernel<<<div_round_up(n_elements, 256), 256, 0, stream>>>(n_elements, decay, grid, grid_tmp);
	NGPB_CUDA_CHECK(cudaGetLastError());
}
void nerf_update_bitfield(cudaStream_t stream, uint32_t max_cascade, const float* grid, uint8_t* bitfield, float* mean_density) {
	NGPB_CUDA_CHECK(cudaMemsetAsync(mean_density, 0, sizeof(float), stream));
	mean_density_kernel<<<sm_count() * 4, 256, 0, stream>>>(kNerfGridCells, grid, mean_density);
	const uint32_t n_bytes = kNerfGridCells / 8;
	grid_to_bitfield_kernel<<<div_round_up(n_bytes * kNerfCascades, 256), 256, 0, stream>>>(n_bytes * kNerfCascades, n_bytes * (max_cascade + 1), grid, bitfield, mean_density);
	for (uint32_t level = 1; level < kNerfCascades; ++level) {
		bitfield_max_pool_kernel<<<div_round_up(kNerfGridCells / 64, 256), 256, 0, stream>>>(kNerfGridCells / 64,
			bitfield + (size_t)grid_mip_offset(level - 1) / 8, bitfield + (size_t)grid_mip_offset(level) / 8);
	}
	NGPB_CUDA_CHECK(cudaGetLastError());
}

} // namespace ngpb
