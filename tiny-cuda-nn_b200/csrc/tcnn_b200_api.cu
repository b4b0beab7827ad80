// libtcnn_b200.so — host side of the C ABI declared in include/tcnn_b200.h.
// Resolves the tcnn JSON config into a plan (level table, MLP shape, loss, Adam hyper-parameters) and launches the
// sm_100a kernels of grid_kernels.cuh / mlp_kernels.cuh / optim_kernels.cuh.  No CPU compute path exists here.
#include "../../include/tcnn_b200.h"

#include <cuda_runtime.h>

#include <algorithm>
#include <cmath>
#include <cstring>
#include <limits>
#include <map>
#include <mutex>
#include <memory>
#include <random>
#include <string>
#include <vector>

#include "grid_kernels.cuh"
#include "json_mini.h"
#include "mlp_kernels.cuh"  // includes grid_kernels.cuh (fused scatter)
#include "mlp_wide_kernels.cuh"
#include "oneblob_kernels.cuh"
#include "optim_kernels.cuh"

using namespace b200;

static thread_local std::string g_last_error;

#define API_BEGIN try {
#define API_END                                                                                   \
	return 0;                                                                                     \
	}                                                                                             \
	catch (const std::exception& e) { g_last_error = e.what(); return 1; }                        \
	catch (...) { g_last_error = "unknown error"; return 1; }

#define FAIL(msg) throw std::runtime_error(std::string(__FILE__) + ":" + std::to_string(__LINE__) + " " + (msg))
#define CHECK(cond, msg) do { if (!(cond)) FAIL(std::string(msg) + " (" #cond ")"); } while (0)
#define CUDA_OK(call) do { cudaError_t e_ = (call); if (e_ != cudaSuccess) FAIL(std::string(#call " failed: ") + cudaGetErrorString(e_)); } while (0)

struct tcnn_b200_plan {
	tcnn_b200_info info{};
	GridParams grid{};
	float log2_per_level_scale = 1.f;
	bool device_scales = false;     // level scales were evaluated by the device (CUDA exp2f) rather than libm
	int n_sms = 148;
	// Adam extras
	int activation = ACT_RELU, output_activation = ACT_NONE; // network.activation / network.output_activation (activation.cuh)
	float non_matrix_lr_factor = 1.f, weight_clip = 0.f;
	bool adabound = false;
	bool optimize_matrix_params = true, optimize_non_matrix_params = true; // adam.h:57-58, 77-85: a false one freezes that parameter class
	int grid_type = 0;              // GRID_HASH / GRID_DENSE / GRID_TILED (GridType, grid_interface.h)
	uint32_t hash_type = 0;         // HashKind (grid_kernels.cuh): "hash" of the encoding config (grid.h:1151), default CoherentPrime
	int enc_kind = 0;               // ENC_GRID / ENC_IDENTITY / ENC_ONEBLOB
	uint32_t oneblob_log2_bins = 0;
	bool f32 = false;               // stand-alone encoding with fp32 parameters / output ("precision": "fp32"; Precision::Fp32 of cpp_api.h)
	int kind = 0;                   // PLAN_FULL: encoding + network; PLAN_NETWORK: Identity encoding + network; PLAN_ENCODING: grid encoding alone
	bool no_warp_specialization = getenv("TCNN_B200_NO_WS") != nullptr; // development switch: fused kernel without the scatter warpgroup
	// optional per-kernel timing (tcnn_b200_profile_*): CUDA events recorded around every launch on the launch stream
	// fork/join for the fused training step: Adam on finished level groups overlaps the scatter of the next group
	cudaStream_t aux = nullptr;
	cudaEvent_t ev[8] = {};
	bool aux_ready = false;
	bool profiling = false;
	struct Span { int kind; cudaEvent_t e0, e1; };
	std::vector<Span> spans;
	// static graphs of tcnn_b200_training_step, one per (batch size, buffers) - see StepGraph below
	uint64_t version = 0;           // bumped by every setter that changes what a captured step would do
	int graph_mode = -1;            // -1: environment TCNN_B200_GRAPH (default on), 0: off, 1: on
	cudaStream_t capture_stream = nullptr;
	std::vector<struct StepGraph*> step_graphs;
	uint64_t graph_clock = 0;
	uint32_t graph_captures = 0, graph_replays = 0, graph_misses_in_a_row = 0;
	bool graph_broken = false;      // capture failed or the Adam nodes were not found: direct launches from then on
	// host-batch pipeline (tcnn_b200_host_batch_begin / _end): two staging slots, copy engines on their own streams
	struct HostPipe {
		bool ready = false;
		cudaStream_t h2d = nullptr, d2h = nullptr;
		cudaEvent_t filled[2] = {}, consumed[2] = {}, loss_written[2] = {}, loss_copied[2] = {};
		const void* holds_x[2] = {nullptr, nullptr};   // host batch a slot was filled from (or is being filled from)
		const void* holds_t[2] = {nullptr, nullptr};
		uint32_t holds_B[2] = {0, 0};
		void* staging[2] = {nullptr, nullptr};
		uint32_t slot = 0;          // slot of the step between begin and end
		bool open = false;
	} pipe;
};

enum PlanKind { PLAN_FULL = 0, PLAN_NETWORK = 1, PLAN_ENCODING = 2 };
enum GridTypeKind { GRID_HASH = 0, GRID_DENSE = 1, GRID_TILED = 2 };
enum EncKind { ENC_GRID = 0, ENC_IDENTITY = 1, ENC_ONEBLOB = 2 };
enum ProfileKind { PK_GRID_FWD = 0, PK_MLP = 1, PK_FINALIZE = 2, PK_GRID_BWD = 3, PK_ADAM = 4, PK_COUNT = 5 };

struct ProfileScope {
	tcnn_b200_plan* p; cudaStream_t s; cudaEvent_t e1{};
	ProfileScope(const tcnn_b200_plan* plan, cudaStream_t stream, int kind) : p(const_cast<tcnn_b200_plan*>(plan)), s(stream) {
		if (!p->profiling) { p = nullptr; return; }
		tcnn_b200_plan::Span sp; sp.kind = kind;
		cudaEventCreate(&sp.e0); cudaEventCreate(&sp.e1);
		cudaEventRecord(sp.e0, s);
		e1 = sp.e1;
		p->spans.push_back(sp);
	}
	~ProfileScope() { if (p) cudaEventRecord(e1, s); }
};

static uint32_t next_multiple_u32(uint32_t v, uint32_t m) { return (v + m - 1) / m * m; }

// Per-level constants of grid_index (common_device.h:690-707) for a given (scale, size): strides with uint32 wrap.
static void fill_level(GridLevel& lv, float scale, uint32_t size, uint32_t offset, uint32_t n_dims, bool hash_type, uint32_t hash_kind) {
	lv.scale = scale;
	lv.resolution = (uint32_t)std::ceil(scale) + 1; // grid_resolution, common_device.h:716-718
	lv.size = size;
	lv.offset = offset;
	uint32_t stride = 1, d = 0;
	for (uint32_t k = 0; k < (uint32_t)MAX_DIMS; ++k) lv.stride[k] = 0;
	for (; d < n_dims && stride <= size; ++d) {
		lv.stride[d] = stride;
		stride *= lv.resolution; // uint32 wrap on purpose
	}
	lv.n_dense_dims = d;
	lv.hashed = (hash_type && size < stride) ? 1u : 0u; // Dense / Tiled grids never hash: dense index with wrap, % size
	lv.replicated = 0; lv.rep_offset = 0;
	lv.magic = fastmod_magic(size ? size : 1u);
	const bool pow2 = size && (size & (size - 1)) == 0;
	lv.mode = lv.hashed ? LEVEL_HASH : pow2 ? LEVEL_DENSE_POW2 : LEVEL_DENSE_MOD; // hashed levels have size 2^log2_hashmap_size
	lv.hash_type = hash_kind;
}

constexpr uint32_t N_REPLICAS = 32;          // private copies of the contended coarse gradient levels
constexpr uint32_t REPLICATE_MAX_ENTRIES = 8192; // levels up to this many entries are replicated ...
constexpr uint32_t REPLICATE_MIN_BATCH = 1u << 14; // ... when the batch is large enough for contention to matter

// Plans the specialised gather / scatter kernels do not cover (a hash other than CoherentPrime, F = 1, a per-sample max level) run the
// general per-(sample, level) kernels: no replicas, no fusion into the MLP kernel.
static bool general_grid(const tcnn_b200_plan* p) {
	return p->enc_kind == 0 /* ENC_GRID */ && p->info.n_levels > 0 && (p->hash_type != HASH_COHERENT_PRIME || p->info.n_features_per_level == 1 || p->grid.max_level_gpu != nullptr);
}
// GridParams for the backward launch: marks the replicated levels (if a replica buffer is available).
static GridParams backward_params(const tcnn_b200_plan* p, uint32_t B, bool have_replicas) {
	GridParams gp = p->grid;
	gp.n_replicas = N_REPLICAS; gp.rep_entries = 0;
	static const bool never = getenv("TCNN_B200_NO_REPLICAS") != nullptr; // development switch (tests/dev/parity_probe.py): plain fp16 reductions on every level
	if (!have_replicas || B < REPLICATE_MIN_BATCH || never || general_grid(p)) return gp;
	for (uint32_t l = 0; l < gp.n_levels; ++l) {
		if (gp.level[l].size <= REPLICATE_MAX_ENTRIES) {
			gp.level[l].replicated = 1;
			gp.level[l].rep_offset = gp.rep_entries;
			gp.rep_entries += gp.level[l].size;
		}
	}
	return gp;
}
static size_t replica_bytes(const tcnn_b200_plan* p) {
	if (p->info.n_features_per_level == 1) return (size_t)p->info.n_params_grid * 4; // F = 1: this region holds the fp32 copy of the table gradient instead
	size_t entries = 0;
	for (uint32_t l = 0; l < p->info.n_levels; ++l) if (p->grid.level[l].size <= REPLICATE_MAX_ENTRIES) entries += p->grid.level[l].size; // all levels, whatever max_level says
	return entries * p->info.n_features_per_level * 2 * N_REPLICAS;
}

static void build_levels(tcnn_b200_plan* p, const float* scales) {
	const tcnn_b200_info& in = p->info;
	for (uint32_t l = 0; l < in.n_levels; ++l) {
		fill_level(p->grid.level[l], scales[l], in.level_offset[l + 1] - in.level_offset[l], in.level_offset[l], in.n_input_dims, p->grid_type == GRID_HASH, p->hash_type);
		p->info.level_scale[l] = scales[l];
		p->info.level_resolution[l] = p->grid.level[l].resolution;
	}
}

// The reference evaluates grid_scale inside its kernels with CUDA's exp2f; do the same so indices are bit-identical
// even for non-power-of-two per_level_scale.  Falls back to libm only when no device is present (CPU-side tests).
static void try_device_scales(tcnn_b200_plan* p) {
	int n_dev = 0;
	if (cudaGetDeviceCount(&n_dev) != cudaSuccess || n_dev == 0) { cudaGetLastError(); return; }
	int dev = 0;
	if (cudaGetDevice(&dev) != cudaSuccess) { cudaGetLastError(); return; }
	cudaDeviceGetAttribute(&p->n_sms, cudaDevAttrMultiProcessorCount, dev);
	float* d_scales = nullptr;
	if (cudaMalloc(&d_scales, sizeof(float) * MAX_LEVELS) != cudaSuccess) { cudaGetLastError(); return; }
	grid_scale_kernel<<<1, MAX_LEVELS>>>(p->info.n_levels, p->log2_per_level_scale, p->info.base_resolution, d_scales);
	float h[MAX_LEVELS];
	const cudaError_t e = cudaMemcpy(h, d_scales, sizeof(float) * p->info.n_levels, cudaMemcpyDeviceToHost);
	cudaFree(d_scales);
	if (e != cudaSuccess) { cudaGetLastError(); return; }
	build_levels(p, h);
	p->device_scales = true;
}

// string_to_activation (src/common_host.cu:74-100).  Sine has no backward in the reference's fused kernels
// (common_device.h:262-266: "requires stored pre-activations, which we don't have") and is not offered.
static int parse_activation(const std::string& name) {
	const std::string n = jsonmini::lower(name);
	if (n == "none") return ACT_NONE;
	if (n == "relu") return ACT_RELU;
	if (n == "leakyrelu") return ACT_LEAKY_RELU;
	if (n == "exponential") return ACT_EXPONENTIAL;
	if (n == "sigmoid") return ACT_SIGMOID;
	if (n == "squareplus") return ACT_SQUAREPLUS;
	if (n == "softplus") return ACT_SOFTPLUS;
	if (n == "tanh") return ACT_TANH;
	FAIL("Invalid activation name: " + name + " (None, ReLU, LeakyReLU, Exponential, Sigmoid, Squareplus, Softplus, Tanh)");
}

static void parse_optimizer(tcnn_b200_plan* p, const jsonmini::Value& o) {
	tcnn_b200_info& in = p->info;
	if (o.contains("otype")) {
		const std::string t = jsonmini::lower(o.string("otype", "adam"));
		CHECK(t == "adam", "optimizer otype '" + t + "' is outside the hot path (only Adam, optimizers/adam.h)");
	}
	// update_hyperparams, optimizers/adam.h:210-258
	in.learning_rate = (float)o.number("learning_rate", in.learning_rate);
	in.beta1 = (float)o.number("beta1", in.beta1);
	in.beta2 = (float)o.number("beta2", in.beta2);
	in.epsilon = (float)o.number("epsilon", in.epsilon);
	in.l2_reg = (float)o.number("l2_reg", in.l2_reg);
	in.relative_decay = (float)o.number("relative_decay", in.relative_decay);
	in.absolute_decay = (float)o.number("absolute_decay", in.absolute_decay);
	p->adabound = o.boolean("adabound", p->adabound);
	p->weight_clip = (float)o.number("clipping_magnitude", p->weight_clip);
	p->non_matrix_lr_factor = (float)o.number("non_matrix_learning_rate_factor", p->non_matrix_lr_factor);
	p->optimize_matrix_params = o.boolean("optimize_matrix_params", p->optimize_matrix_params);
	p->optimize_non_matrix_params = o.boolean("optimize_non_matrix_params", p->optimize_non_matrix_params);
	// anything else would be silently ignored: a misspelt key must not train with defaults
	static const char* known[] = {"otype", "learning_rate", "beta1", "beta2", "epsilon", "l2_reg", "relative_decay", "absolute_decay", "adabound",
		"clipping_magnitude", "non_matrix_learning_rate_factor", "optimize_matrix_params", "optimize_non_matrix_params"};
	for (const std::string& key : o.keys()) {
		bool ok = false;
		for (const char* k : known) ok |= key == k;
		CHECK(ok, "Adam: unknown optimizer key '" + key + "' (optimizers/adam.h:210-258)");
	}
}

// ---- plan ------------------------------------------------------------------------------------------
extern "C" const char* tcnn_b200_last_error(void) { return g_last_error.c_str(); }
extern "C" uint32_t tcnn_b200_abi_version(void) { return 1; }
extern "C" const char* tcnn_b200_arch(void) { return "sm_100a"; }
extern "C" uint32_t tcnn_b200_batch_size_granularity(void) { return 256; }

extern "C" int tcnn_b200_plan_create(const char* json_utf8, uint32_t n_input_dims, uint32_t n_output_dims, tcnn_b200_plan** out) {
	API_BEGIN
	CHECK(json_utf8 && out, "null argument");
	const jsonmini::Value cfg = jsonmini::parse(json_utf8);
	const jsonmini::Value enc = cfg.object("encoding"), net = cfg.object("network"), loss = cfg.object("loss"), opt = cfg.object("optimizer");
	std::unique_ptr<tcnn_b200_plan> p(new tcnn_b200_plan());
	tcnn_b200_info& in = p->info;
	in.n_input_dims = n_input_dims;
	in.n_output_dims = n_output_dims;

	// Which reference object the config describes (src/cpp_api.cu:143-165): encoding + network = NetworkWithInputEncoding,
	// network alone = create_network (Identity encoding in front), encoding alone = create_encoding (no padding to 16).
	const bool has_encoding = cfg.contains("encoding"), has_network = cfg.contains("network");
	CHECK(has_encoding || has_network, "config needs an \"encoding\" and/or a \"network\" object");
	p->kind = has_encoding ? (has_network ? PLAN_FULL : PLAN_ENCODING) : PLAN_NETWORK;
	// create_encoding(n_input_dims, encoding, Precision) of cpp_api.h:71-75,115: a stand-alone encoding may be asked for in fp32
	const std::string precision = jsonmini::lower(cfg.string("precision", "fp16"));
	CHECK(precision == "fp16" || precision == "fp32", "precision must be \"fp16\" or \"fp32\"");
	p->f32 = precision == "fp32";
	CHECK(!p->f32 || p->kind == PLAN_ENCODING, "fp32 precision exists for a stand-alone encoding only (networks are fp16, TCNN_HALF_PRECISION)");
	uint32_t n_features = 0;
	if (has_encoding) {
	// --- encoding: create_grid_encoding (encodings/grid.h:1143-1208) ---
	const std::string etype = jsonmini::lower(enc.string("otype", "Grid"));
	if (etype == "oneblob") {
		// OneBlobEncoding (encodings/oneblob.h:162-175; created by src/encoding.cu with n_bins default 16)
		CHECK(has_network, "a stand-alone OneBlob encoding is outside the hot path (OneBlob feeds the fused MLP, config #2)");
		const uint32_t n_bins = (uint32_t)enc.number("n_bins", 16);
		CHECK(n_bins >= 2 && (n_bins & (n_bins - 1)) == 0, "Number of bins must be a power of 2");
		CHECK(n_input_dims >= 1 && (uint64_t)n_input_dims * n_bins <= 128, "OneBlob: n_input_dims * n_bins must be <= 128 here");
		p->enc_kind = ENC_ONEBLOB;
		while ((1u << p->oneblob_log2_bins) < n_bins) ++p->oneblob_log2_bins;
		n_features = n_input_dims * n_bins;
		in.encoded_width = next_multiple_u32(n_features, 16u);
		p->grid.n_features = n_features;
		p->grid.padded_width = in.encoded_width;
		try_device_scales(p.get());
	} else {
	CHECK(etype == "hashgrid" || etype == "grid" || etype == "densegrid" || etype == "tiledgrid",
		"encoding otype '" + etype + "' is outside the hot path (HashGrid / DenseGrid / TiledGrid and OneBlob; SURVEY.md section 8)");
	// create_grid_encoding (grid.h:1180-1208): DenseGrid / TiledGrid fix the type, otherwise "type" decides (default Hash)
	const std::string gtype = etype == "densegrid" ? "dense" : etype == "tiledgrid" ? "tiled" : jsonmini::lower(enc.string("type", "Hash"));
	if (gtype == "hash") p->grid_type = GRID_HASH;
	else if (gtype == "dense") p->grid_type = GRID_DENSE;
	else if (gtype == "tiled") p->grid_type = GRID_TILED;
	else FAIL("GridEncoding: invalid grid type '" + gtype + "'");
	{ // string_to_hash_type (grid.h:1106-1119)
		const std::string h = jsonmini::lower(enc.string("hash", "CoherentPrime"));
		if (h == "coherentprime") p->hash_type = HASH_COHERENT_PRIME;
		else if (h == "prime") p->hash_type = HASH_PRIME;
		else if (h == "reversedprime") p->hash_type = HASH_REVERSED_PRIME;
		else if (h == "rng") p->hash_type = HASH_RNG;
		else FAIL("Invalid hash type: " + h);
	}
	const std::string interp = jsonmini::lower(enc.string("interpolation", "Linear"));
	CHECK(interp == "nearest" || interp == "linear" || interp == "smoothstep", "interpolation '" + interp + "' does not exist (Nearest, Linear, Smoothstep)");
	p->grid.smoothstep = interp == "smoothstep" ? 1u : 0u;
	p->grid.nearest = interp == "nearest" ? 1u : 0u;
	p->grid.stochastic = enc.boolean("stochastic_interpolation", false) ? 1u : 0u;
	CHECK(!(p->f32 && p->grid.stochastic), "stochastic_interpolation is implemented for fp16 precision");
	{ // pcg32{1337} (pcg32.h:54-60 seed(initstate, initseq = 1)): the generator random_val() of the scatter starts from
		const uint64_t mult = 0x5851f42d4c957f2dULL, inc = 3u;
		uint64_t state = 0u;
		state = state * mult + inc;
		state += 1337u;
		state = state * mult + inc;
		p->grid.rng_state = state; p->grid.rng_inc = inc;
	}
	in.n_features_per_level = (uint32_t)enc.number("n_features_per_level", 2);
	CHECK(in.n_features_per_level == 1 || in.n_features_per_level == 2 || in.n_features_per_level == 4 || in.n_features_per_level == 8,
		"GridEncoding: n_features_per_level must be 1, 2, 4 or 8");
	CHECK(!(in.n_features_per_level == 1 && p->f32), "n_features_per_level = 1 is implemented for fp16 precision");
	if (enc.contains("n_features") || enc.contains("n_grid_features")) {
		CHECK(!enc.contains("n_levels"), "GridEncoding: may not specify n_features and n_levels simultaneously (one determines the other)");
		n_features = (uint32_t)enc.number(enc.contains("n_features") ? "n_features" : "n_grid_features", 0);
	} else {
		n_features = in.n_features_per_level * (uint32_t)enc.number("n_levels", 16);
	}
	CHECK(n_features % in.n_features_per_level == 0, "GridEncoding: n_features must be a multiple of n_features_per_level");
	in.n_levels = n_features / in.n_features_per_level;
	CHECK(in.n_levels >= 1 && in.n_levels <= (uint32_t)MAX_LEVELS, "GridEncoding: n_levels must be in [1, 32] here");
	in.log2_hashmap_size = (uint32_t)enc.number("log2_hashmap_size", 19);
	CHECK(in.log2_hashmap_size >= 3 && in.log2_hashmap_size <= 28, "log2_hashmap_size out of range");
	in.base_resolution = (uint32_t)enc.number("base_resolution", 16);
	in.per_level_scale = (float)enc.number("per_level_scale", 2.0);
	CHECK(n_input_dims >= 2 && n_input_dims <= 4, "GridEncoding: number of input dims must be 2, 3 or 4");
	// set_alignment(16) in front of FullyFusedMLP (network_with_input_encoding.h:47); a stand-alone encoding is not padded by
	// the reference (alignment 0, src/cpp_api.cu:156-165) - its rows are stored here at the next multiple of 8 columns
	in.encoded_width = next_multiple_u32(n_features, has_network ? 16u : 8u);
	p->log2_per_level_scale = std::log2(in.per_level_scale);

	// offset table on the host, as the reference's constructor does (grid.h:692-716)
	float host_scales[MAX_LEVELS];
	uint32_t offset = 0;
	for (uint32_t l = 0; l < in.n_levels; ++l) {
		host_scales[l] = exp2f(l * p->log2_per_level_scale) * in.base_resolution - 1.0f;
		const uint32_t resolution = (uint32_t)std::ceil(host_scales[l]) + 1;
		const uint32_t max_params = std::numeric_limits<uint32_t>::max() / 2;
		uint32_t n = max_params;
		if (!(std::pow((float)resolution, (float)n_input_dims) > (float)max_params)) {
			n = 1;
			for (uint32_t d = 0; d < n_input_dims; ++d) n *= resolution;
		}
		n = next_multiple_u32(n, 8u);
		if (p->grid_type == GRID_HASH) n = std::min(n, 1u << in.log2_hashmap_size);   // grid.h:702-712
		else if (p->grid_type == GRID_TILED) {
			uint32_t tile = 1;
			for (uint32_t d = 0; d < n_input_dims; ++d) tile *= in.base_resolution;     // powi(base_resolution, N_POS_DIMS)
			n = std::min(n, tile);
			// every kernel reads / reduces aligned 16-byte slots of the level tables: level offsets must stay multiples of 8 entries.
			// The reference accepts base_resolution^D that is not (e.g. 5 or 6 in 2-D); here that is an error, not a misaligned access.
			CHECK(n % 8 == 0, "TiledGrid: base_resolution^n_input_dims (" + std::to_string(tile) + ") must be a multiple of 8 in this build (level tables are accessed in aligned 16-byte slots)");
		}
		in.level_offset[l] = offset;
		CHECK((uint64_t)offset + n < (1ull << 32) / in.n_features_per_level, "GridEncoding: more than 2^32 grid parameters");
		CHECK((uint64_t)n * in.n_features_per_level * 4 <= 0xFFFFFFFFull, "GridEncoding: a single level of more than 4 GiB (fp32) is not supported (the kernels address entries inside a level with 32-bit byte offsets)");
		offset += n;
	}
	in.level_offset[in.n_levels] = offset;
	in.n_params_grid = (uint64_t)offset * in.n_features_per_level;
	p->grid.n_levels = in.n_levels;
	p->grid.n_levels_total = in.n_levels;
	p->grid.n_features = n_features;
	p->grid.padded_width = in.encoded_width;
	build_levels(p.get(), host_scales);
	try_device_scales(p.get());
	}
	} else {
		p->enc_kind = ENC_IDENTITY;
		// Identity encoding (encodings/identity.h:46-90): fp32 -> fp16 copy, padded to a multiple of 16 columns with ones
		CHECK(n_input_dims >= 1 && n_input_dims <= 64, "Network: n_input_dims must be in [1, 64] here");
		in.encoded_width = next_multiple_u32(n_input_dims, 16);
		try_device_scales(p.get());
	}

	if (has_network) {
	// --- network: create_network (src/network.cu:97-140), FullyFusedMLP ctor (src/fully_fused_mlp.cu:635-678) ---
	const std::string ntype = jsonmini::lower(net.string("otype", "FullyFusedMLP"));
	CHECK(ntype == "fullyfusedmlp" || ntype == "megakernelmlp", "network otype '" + ntype + "' is outside the hot path (FullyFusedMLP only)");
	in.n_neurons = (uint32_t)net.number("n_neurons", 128);
	in.n_hidden_layers = (uint32_t)net.number("n_hidden_layers", 5);
	CHECK(in.n_neurons == 16 || in.n_neurons == 32 || in.n_neurons == 64 || in.n_neurons == 128, "FullyFusedMLP only supports 16, 32, 64, and 128 neurons, but got " + std::to_string(in.n_neurons)); // fully_fused_mlp.cu:862-870
	if (in.n_neurons <= 64) CHECK(in.n_hidden_layers >= 1 && in.n_hidden_layers <= (uint32_t)MLP_MAX_HIDDEN, "FullyFusedMLP: n_hidden_layers must be in [1, 5] at up to 64 neurons");
	else CHECK(in.n_hidden_layers >= 1 && in.n_hidden_layers <= (uint32_t)WIDE_MAX_HIDDEN, "FullyFusedMLP: n_hidden_layers must be in [1, 8] at 128 neurons");
	p->activation = parse_activation(net.string("activation", "ReLU"));               // defaults: src/network.cu:107-113
	p->output_activation = parse_activation(net.string("output_activation", "None"));
	CHECK(n_output_dims >= 1 && n_output_dims <= 16, "n_output_dims must be in [1, 16] (fused last layer, fully_fused_mlp.cu:544)");
	CHECK(in.encoded_width <= 128, "encoded width above 128 is not implemented");
	if (in.n_neurons == 128)
		CHECK(wide_fwd_smem(in.n_hidden_layers, in.encoded_width, 1) <= 232448u, "FullyFusedMLP 128: the weight matrices of this many hidden layers do not fit shared memory");
	in.padded_output_width = 16;
	in.n_params_mlp = (uint64_t)in.n_neurons * in.encoded_width + (uint64_t)(in.n_hidden_layers - 1) * in.n_neurons * in.n_neurons + (uint64_t)16 * in.n_neurons;
	} else {
		in.n_output_dims = n_features;             // Encoding.n_output_dims (modules.py:272-275)
		in.padded_output_width = in.encoded_width;
		in.n_params_mlp = 0;
	}
	in.n_params = in.n_params_mlp + in.n_params_grid;
	CHECK(in.n_params < (1ull << 32), "more than 2^32 parameters");

	// --- loss: create_loss (src/loss.cu:53-93; default RelativeL2) ---
	const std::string ltype = jsonmini::lower(loss.string("otype", "RelativeL2"));
	if (ltype == "l2") in.loss_type = LOSS_L2;
	else if (ltype == "relativel2") in.loss_type = LOSS_RELATIVE_L2;
	else FAIL("loss otype '" + ltype + "' is outside the hot path (L2, RelativeL2)");

	// --- optimizer: Adam defaults (optimizers/adam.h:312-326) ---
	in.learning_rate = 1e-3f; in.beta1 = 0.9f; in.beta2 = 0.999f; in.epsilon = 1e-8f; in.l2_reg = 1e-8f;
	in.relative_decay = 0.f; in.absolute_decay = 0.f;
	parse_optimizer(p.get(), opt);

	*out = p.release();
	API_END
}

struct StepGraph;
static void drop_step_graphs(tcnn_b200_plan* p);
extern "C" void tcnn_b200_plan_destroy(tcnn_b200_plan* plan) {
	if (plan) {
		drop_step_graphs(plan);
		if (plan->capture_stream) cudaStreamDestroy(plan->capture_stream);
		if (plan->pipe.ready) {
			cudaStreamDestroy(plan->pipe.h2d); cudaStreamDestroy(plan->pipe.d2h);
			for (int k = 0; k < 2; ++k) { cudaEventDestroy(plan->pipe.filled[k]); cudaEventDestroy(plan->pipe.consumed[k]); cudaEventDestroy(plan->pipe.loss_written[k]); cudaEventDestroy(plan->pipe.loss_copied[k]); }
		}
	}
	if (plan && plan->aux_ready) {
		cudaStreamDestroy(plan->aux);
		for (auto& e : plan->ev) cudaEventDestroy(e);
	}
	delete plan;
}
static void ensure_aux(const tcnn_b200_plan* plan_) {
	tcnn_b200_plan* p = const_cast<tcnn_b200_plan*>(plan_);
	if (p->aux_ready) return;
	CUDA_OK(cudaStreamCreateWithFlags(&p->aux, cudaStreamNonBlocking));
	for (auto& e : p->ev) CUDA_OK(cudaEventCreateWithFlags(&e, cudaEventDisableTiming));
	p->aux_ready = true;
}

// GridEncoding::set_max_level (grid_interface.h:101-107).  Levels with level >= max_level * n_levels + 1e-3 (fp32, exactly the
// comparison of grid.h:69-75) encode to zero and get no gradient (grid.h:236-243).  The kernels stop at GridParams::n_levels
// and leave the columns behind it zero, so the setter only moves that bound; parameter layout and counts do not change.
extern "C" int tcnn_b200_plan_set_max_level(tcnn_b200_plan* plan, float max_level) {
	API_BEGIN
	CHECK(plan, "null argument");
	CHECK(plan->enc_kind == ENC_GRID && plan->info.n_levels > 0, "max_level belongs to grid encodings");
	const uint32_t L = plan->info.n_levels, F = plan->info.n_features_per_level;
	const float scaled = (max_level * (float)(L * F)) / (float)F;
	uint32_t active = 0;
	for (uint32_t l = 0; l < L; ++l) if (!((float)l >= scaled + 1e-3f)) active = l + 1;
	plan->grid.n_levels = active;
	plan->grid.n_features = active * F;
	++plan->version;
	API_END
}

// GridEncoding::set_max_level_gpu (grid_interface.h:113-115): a device array of one max level per sample of the batches to come
// (nullptr switches it off).  The plan then runs the general grid kernels (grid_general_forward_kernel ...).
extern "C" int tcnn_b200_plan_set_max_level_gpu(tcnn_b200_plan* plan, const float* max_level_per_sample) {
	API_BEGIN
	CHECK(plan, "null argument");
	CHECK(plan->enc_kind == ENC_GRID && plan->info.n_levels > 0, "max_level belongs to grid encodings");
	CHECK(!plan->f32 || !max_level_per_sample, "the per-sample max level is implemented for fp16 precision");
	plan->grid.max_level_gpu = max_level_per_sample;
	++plan->version;
	API_END
}

extern "C" int tcnn_b200_plan_info(const tcnn_b200_plan* plan, tcnn_b200_info* info) {
	API_BEGIN
	CHECK(plan && info, "null argument");
	*info = plan->info;
	API_END
}

extern "C" int tcnn_b200_plan_update_optimizer(tcnn_b200_plan* plan, const char* json_utf8) {
	API_BEGIN
	CHECK(plan && json_utf8, "null argument");
	parse_optimizer(plan, jsonmini::parse(json_utf8));
	++plan->version;
	API_END
}

// workspace layout: [encoded B x E fp16][dL_dencoded B x E fp16][dW partials n_ctas x P_mlp fp32][loss partials n_ctas fp32]
struct Workspace {
	__half* encoded; __half* dL_denc; float* dw_partials; float* loss_partials; __half* replicas;
	// 128-neuron path only: hidden activations and pre-activation gradients [H][B][128], dL/d(output) [B][16]
	__half* wide_hidden; __half* wide_gtmp; __half* wide_dy;
};
static bool is_wide(const tcnn_b200_plan* p) { return p->info.n_neurons == 128; }
// 64 neurons: CTAs of the persistent MLP kernel = number of weight-gradient / loss partials.
// 128 neurons: number of batch splits of the weight-gradient kernel (one CTA per split and matrix, one CTA per SM).
static uint32_t mlp_grid_size(const tcnn_b200_plan* p, uint32_t n_tiles, bool train) {
	if (is_wide(p) && train) return std::max(1u, std::min(n_tiles, (uint32_t)p->n_sms / (p->info.n_hidden_layers + 1))); // one CTA per SM (192 KB ring)
	// resident CTAs per SM of the 64-neuron kernel: 2 when training (100 KB, 256 TMEM columns each), 3 for inference; the
	// 5-hidden-layer / wide-input shapes need 512 TMEM columns or > 113 KB: one
	const bool big = p->info.n_hidden_layers > 4 || p->info.encoded_width > 64;
	const uint32_t per_sm = train ? (big ? 1u : 2u) : (big ? 2u : 3u);
	return std::max(1u, std::min(n_tiles, (uint32_t)p->n_sms * per_sm));
}
// warpgroups (independent tiles) per CTA of the 128-neuron forward / backward kernels: two if shared memory allows
static uint32_t wide_nwg_fwd(const tcnn_b200_plan* p) { return wide_fwd_smem(p->info.n_hidden_layers, p->info.encoded_width, 2) <= 232448u ? 2u : 1u; }
static uint32_t wide_nwg_bwd(const tcnn_b200_plan* p, bool want_dx) { return wide_bwd_smem(p->info.n_hidden_layers, p->info.encoded_width, 2, want_dx) <= 232448u ? 2u : 1u; }
static uint32_t wide_fwd_ctas(const tcnn_b200_plan* p, uint32_t n_tiles) { const uint32_t g = wide_nwg_fwd(p); return std::max(1u, std::min((n_tiles + g - 1) / g, (uint32_t)p->n_sms)); }
// number of loss partials the training forward leaves (the 64-neuron kernel: one per CTA = per weight-gradient partial)
static uint32_t mlp_loss_parts(const tcnn_b200_plan* p, uint32_t n_tiles, uint32_t n_parts) {
	return is_wide(p) ? wide_fwd_ctas(p, n_tiles) * wide_nwg_fwd(p) : n_parts;
}
static size_t align256(size_t v) { return (v + 255) / 256 * 256; }
// encoded batch: rows [B][E] for the 64-neuron kernel, tile images (whole 64-column sub-tiles) for the 128-neuron kernels
static uint32_t image_subs(const tcnn_b200_plan* p) { return is_wide(p) ? wide_in_sub(p->info.encoded_width) : 0u; }
static size_t encoded_bytes(const tcnn_b200_plan* p, uint32_t B) {
	return align256((size_t)B * (is_wide(p) ? wide_in_sub(p->info.encoded_width) * 64u : p->info.encoded_width) * 2);
}
static Workspace carve(const tcnn_b200_plan* p, uint32_t B, void* ws) {
	Workspace w;
	uint8_t* b = (uint8_t*)ws;
	const size_t act = encoded_bytes(p, B);
	const uint32_t max_ctas = (uint32_t)p->n_sms * 4;
	w.encoded = (__half*)b; b += act;
	w.dL_denc = (__half*)b; b += act;
	w.dw_partials = (float*)b; b += align256((size_t)max_ctas * p->info.n_params_mlp * 4);
	w.loss_partials = (float*)b; b += align256((size_t)max_ctas * 4);
	w.replicas = (__half*)b; b += align256(replica_bytes(p));
	w.wide_hidden = w.wide_gtmp = w.wide_dy = nullptr;
	if (is_wide(p)) {
		const size_t layer = align256((size_t)p->info.n_hidden_layers * B * WIDE_W * 2);
		w.wide_hidden = (__half*)b; b += layer;
		w.wide_gtmp = (__half*)b; b += layer;
		w.wide_dy = (__half*)b;
	}
	return w;
}
extern "C" size_t tcnn_b200_workspace_bytes(const tcnn_b200_plan* plan, uint32_t B) {
	if (!plan) return 0;
	const size_t act = encoded_bytes(plan, B);
	const uint32_t max_ctas = (uint32_t)plan->n_sms * 4;
	const size_t wide = is_wide(plan) ? 2 * align256((size_t)plan->info.n_hidden_layers * B * WIDE_W * 2) + align256((size_t)B * MLP_OUT * 2) : 0;
	return 2 * act + align256((size_t)max_ctas * plan->info.n_params_mlp * 4) + align256((size_t)max_ctas * 4) + align256(replica_bytes(plan)) + wide + 256;
}

static void check_batch(uint32_t B) {
	CHECK(B > 0 && B % 256 == 0, "batch size must be a multiple of 256 (BATCH_SIZE_GRANULARITY, common.h:235), got " + std::to_string(B));
}
static void require_device(const tcnn_b200_plan* p) {
	CHECK(p, "null plan");
	if (!p->device_scales) {
		try_device_scales(const_cast<tcnn_b200_plan*>(p));
		CHECK(p->device_scales, "no CUDA device: libtcnn_b200 has no CPU path");
	}
}
static void launch_check(const char* what) {
	const cudaError_t e = cudaGetLastError();
	if (e != cudaSuccess) FAIL(std::string(what) + " launch failed: " + cudaGetErrorString(e));
}

// ---- init -------------------------------------------------------------------------------------------
extern "C" int tcnn_b200_init_params(const tcnn_b200_plan* plan, void* stream_, uint64_t seed, int trainer_seeding, float* params_fp32) {
	API_BEGIN
	require_device(plan);
	CHECK(params_fp32, "null params");
	cudaStream_t stream = (cudaStream_t)stream_;
	const tcnn_b200_info& in = plan->info;
	uint64_t initstate = seed;
	if (trainer_seeding) { // trainer.h:52-55
		std::seed_seq seq{(uint32_t)seed};
		uint32_t seeds[2];
		seq.generate(seeds, seeds + 2);
		initstate = seeds[0];
	}
	// pcg32::seed(initstate, 1)  (pcg32.h:54-60)
	const uint64_t mult = 0x5851f42d4c957f2dULL;
	uint64_t state = 0, inc = (1ull << 1) | 1ull;
	state = state * mult + inc;
	state += initstate;
	state = state * mult + inc;
	uint64_t base = 0;
	float* dst = params_fp32;
	const uint32_t W = in.n_neurons;
	const uint32_t n_mats = in.n_params_mlp ? in.n_hidden_layers + 1 : 0; // a stand-alone encoding draws its table first
	for (uint32_t k = 0; k < n_mats; ++k) {
		const uint32_t rows = (k + 1 == n_mats) ? in.padded_output_width : W;
		const uint32_t cols = (k == 0) ? in.encoded_width : W;
		const uint32_t n = rows * cols;
		const float scale = std::sqrt(6.0f / (float)(cols + rows)); // gpu_matrix.h:289 (fan_in = cols, fan_out = rows)
		init_xavier_kernel<<<(n + 255) / 256, 256, 0, stream>>>(n, state, inc, base, scale, dst);
		launch_check("init_xavier_kernel");
		base += n; dst += n;
	}
	const uint32_t ng = (uint32_t)in.n_params_grid;
	const uint32_t n_threads = next_multiple_u32((ng + 3) / 4, 128); // random.h:58-63 + n_blocks_linear (128 threads)
	if (ng) {
		init_uniform_kernel<<<(ng + 255) / 256, 256, 0, stream>>>(ng, n_threads, state, inc, base, -1e-4f, 1e-4f, dst);
		launch_check("init_uniform_kernel");
	}
	API_END
}

extern "C" int tcnn_b200_cast_params(void* stream_, uint64_t n, const float* params_fp32, void* params_fp16) {
	API_BEGIN
	CHECK(params_fp32 && params_fp16, "null argument");
	cast_f32_to_f16_kernel<<<(unsigned)((n + 255) / 256), 256, 0, (cudaStream_t)stream_>>>(n, params_fp32, (__half*)params_fp16);
	launch_check("cast_f32_to_f16_kernel");
	API_END
}

// generate_random_uniform<float> (random.h:39-71) for the hosts of the C++ facade (samples draw their training batches with it):
// element i + n_threads * j is the j-th draw of the generator advanced by 4 i, n_threads = 128 * ceil(ceil(n / 4) / 128).
extern "C" int tcnn_b200_generate_random_uniform(void* stream, uint64_t rng_state, uint64_t rng_inc, uint64_t n, float* out, float lower, float upper) {
	API_BEGIN
	CHECK(out || n == 0, "null output");
	CHECK(n < (1ull << 32), "at most 2^32 - 1 elements per call");
	if (n == 0) return 0;
	const uint32_t n32 = (uint32_t)n, n_threads = next_multiple_u32((n32 + 3) / 4, 128);
	init_uniform_kernel<<<(n32 + 255) / 256, 256, 0, (cudaStream_t)stream>>>(n32, n_threads, rng_state, rng_inc, 0, lower, upper, out);
	launch_check("init_uniform_kernel");
	API_END
}

// ---- grid encoding ------------------------------------------------------------------------------------
// NOTE (measured, r01): the gather kernel depends on a large L1 (adjacent corners and the coarse levels hit in L1);
// forcing the max-shared carve-out so that it could co-reside with the MLP kernel doubled its time (48 -> 102 us).
#define GRID_LAUNCH(KERNEL, DD, FF, grid_dim, stream, ...)                                                        \
	do {                                                                                                          \
		KERNEL<DD, FF><<<grid_dim, 256, 0, stream>>>(__VA_ARGS__);                                                \
	} while (0)
#define GRID_DISPATCH(KERNEL, plan, grid_dim, stream, ...)                                                        \
	do {                                                                                                          \
		const uint32_t D_ = (plan)->info.n_input_dims, F_ = (plan)->info.n_features_per_level;                   \
		if (D_ == 2 && F_ == 2) GRID_LAUNCH(KERNEL, 2, 2, grid_dim, stream, __VA_ARGS__);                  \
		else if (D_ == 2 && F_ == 4) GRID_LAUNCH(KERNEL, 2, 4, grid_dim, stream, __VA_ARGS__);             \
		else if (D_ == 2 && F_ == 8) GRID_LAUNCH(KERNEL, 2, 8, grid_dim, stream, __VA_ARGS__);             \
		else if (D_ == 3 && F_ == 2) GRID_LAUNCH(KERNEL, 3, 2, grid_dim, stream, __VA_ARGS__);             \
		else if (D_ == 3 && F_ == 4) GRID_LAUNCH(KERNEL, 3, 4, grid_dim, stream, __VA_ARGS__);             \
		else if (D_ == 3 && F_ == 8) GRID_LAUNCH(KERNEL, 3, 8, grid_dim, stream, __VA_ARGS__);             \
		else if (D_ == 4 && F_ == 2) GRID_LAUNCH(KERNEL, 4, 2, grid_dim, stream, __VA_ARGS__);             \
		else if (D_ == 4 && F_ == 4) GRID_LAUNCH(KERNEL, 4, 4, grid_dim, stream, __VA_ARGS__);             \
		else FAIL("unsupported (n_input_dims, n_features_per_level)");                                            \
	} while (0)

#define GRID_DISPATCH_F1(KERNEL, plan, grid_dim, stream, ...)                                                     \
	do {                                                                                                          \
		const uint32_t D1_ = (plan)->info.n_input_dims;                                                           \
		if ((plan)->info.n_features_per_level != 1) GRID_DISPATCH(KERNEL, plan, grid_dim, stream, __VA_ARGS__);   \
		else if (D1_ == 2) GRID_LAUNCH(KERNEL, 2, 1, grid_dim, stream, __VA_ARGS__);                        \
		else if (D1_ == 3) GRID_LAUNCH(KERNEL, 3, 1, grid_dim, stream, __VA_ARGS__);                        \
		else if (D1_ == 4) GRID_LAUNCH(KERNEL, 4, 1, grid_dim, stream, __VA_ARGS__);                        \
		else FAIL("unsupported n_input_dims");                                                                    \
	} while (0)

#define GRID_DISPATCH_T(KERNEL, TT, plan, grid_dim, stream, ...)                                                  \
	do {                                                                                                          \
		const uint32_t D_ = (plan)->info.n_input_dims, F_ = (plan)->info.n_features_per_level;                   \
		if (D_ == 2 && F_ == 2) KERNEL<2, 2, TT><<<grid_dim, 256, 0, stream>>>(__VA_ARGS__);                     \
		else if (D_ == 2 && F_ == 4) KERNEL<2, 4, TT><<<grid_dim, 256, 0, stream>>>(__VA_ARGS__);                \
		else if (D_ == 2 && F_ == 8) KERNEL<2, 8, TT><<<grid_dim, 256, 0, stream>>>(__VA_ARGS__);                \
		else if (D_ == 3 && F_ == 2) KERNEL<3, 2, TT><<<grid_dim, 256, 0, stream>>>(__VA_ARGS__);                \
		else if (D_ == 3 && F_ == 4) KERNEL<3, 4, TT><<<grid_dim, 256, 0, stream>>>(__VA_ARGS__);                \
		else if (D_ == 3 && F_ == 8) KERNEL<3, 8, TT><<<grid_dim, 256, 0, stream>>>(__VA_ARGS__);                \
		else if (D_ == 4 && F_ == 2) KERNEL<4, 2, TT><<<grid_dim, 256, 0, stream>>>(__VA_ARGS__);                \
		else if (D_ == 4 && F_ == 4) KERNEL<4, 4, TT><<<grid_dim, 256, 0, stream>>>(__VA_ARGS__);                \
		else FAIL("unsupported (n_input_dims, n_features_per_level)");                                            \
	} while (0)

static void encode_forward(const tcnn_b200_plan* plan, cudaStream_t stream, uint32_t B, const float* x, const __half* grid, __half* encoded,
	void* zero_on_the_side = nullptr, size_t zero_bytes = 0, bool image = false) {
	ProfileScope prof(plan, stream, PK_GRID_FWD);
	const dim3 g((B + 255) / 256, plan->info.encoded_width / 8, 1);
	uint4* z16 = zero_bytes ? (uint4*)zero_on_the_side : nullptr;
	const uint32_t nz16 = (uint32_t)(zero_bytes / 16);
	GridParams gp = plan->grid;
	gp.image_subs = image ? image_subs(plan) : 0u;
	if (general_grid(plan)) {
		if (z16) CUDA_OK(cudaMemsetAsync(z16, 0, zero_bytes, stream));
		const dim3 gl((B + 255) / 256, plan->info.encoded_width / plan->info.n_features_per_level, 1);
		GRID_DISPATCH_F1(grid_general_forward_kernel, plan, gl, stream, gp, B, x, grid, encoded);
		launch_check("grid_general_forward_kernel");
		return;
	}
	if (gp.nearest) {
		if (z16) CUDA_OK(cudaMemsetAsync(z16, 0, zero_bytes, stream));
		GRID_DISPATCH(grid_nearest_forward_kernel, plan, g, stream, gp, B, x, grid, encoded);
		launch_check("grid_nearest_forward_kernel");
		return;
	}
	GRID_DISPATCH(grid_forward_kernel, plan, g, stream, gp, B, x, grid, encoded, z16, nz16);
	launch_check("grid_forward_kernel");
}
// backward in three pieces so that several batch chunks can scatter concurrently between one prepare and one finish
static size_t replica_bytes_of(const tcnn_b200_plan* plan, const GridParams& gp) { return (size_t)gp.rep_entries * plan->info.n_features_per_level * 2 * gp.n_replicas; }
static void backward_prepare(const tcnn_b200_plan* plan, cudaStream_t stream, const GridParams& gp, __half* grid_grads, int zero_first, __half* replicas) {
	ProfileScope prof(plan, stream, PK_GRID_BWD);
	if (zero_first) CUDA_OK(cudaMemsetAsync(grid_grads, 0, plan->info.n_params_grid * 2, stream));
	if (plan->info.n_features_per_level == 1) { // fp32 copy of the table gradient (grid.h:846-859)
		CHECK(replicas != nullptr, "n_features_per_level = 1 accumulates its gradients in fp32: a workspace is required");
		CUDA_OK(cudaMemsetAsync(replicas, 0, plan->info.n_params_grid * 4, stream));
	}
	if (gp.rep_entries) CUDA_OK(cudaMemsetAsync(replicas, 0, (size_t)gp.rep_entries * plan->info.n_features_per_level * 2 * gp.n_replicas, stream));
}
static uint32_t n_grad_chunks(const tcnn_b200_plan* plan) { return (plan->grid.n_features + 7) / 8; } // padding chunks carry no gradient
static void backward_scatter(const tcnn_b200_plan* plan, cudaStream_t stream, const GridParams& gp, uint32_t B, const float* x, const __half* dL_denc, __half* grid_grads, __half* replicas,
	uint32_t chunk_begin = 0, uint32_t chunk_end = 0xFFFFFFFFu) {
	ProfileScope prof(plan, stream, PK_GRID_BWD);
	chunk_end = std::min(chunk_end, n_grad_chunks(plan));
	if (chunk_begin >= chunk_end) return;
	const dim3 g((B + 255) / 256, chunk_end - chunk_begin, 1);
	if (general_grid(plan)) { // all levels in one launch, with the first chunk group
		if (chunk_begin != 0) return;
		const dim3 gl((B + 255) / 256, std::max(gp.n_levels, 1u), 1);
		GRID_DISPATCH_F1(grid_general_backward_kernel, plan, gl, stream, gp, B, x, dL_denc, grid_grads, (float*)replicas);
		launch_check("grid_general_backward_kernel");
		return;
	}
	if (gp.stochastic && !gp.nearest) { // one random corner per level and sample (grid.h:284-303); Nearest returns before it there too
		GRID_DISPATCH(grid_stochastic_backward_kernel, plan, g, stream, gp, B, x, dL_denc, grid_grads, chunk_begin);
		launch_check("grid_stochastic_backward_kernel");
		return;
	}
	if (gp.nearest) { // one RED per level and sample straight into the gradient buffer; the (zeroed) replicas stay zero
		GRID_DISPATCH(grid_nearest_backward_kernel, plan, g, stream, gp, B, x, dL_denc, grid_grads, chunk_begin);
		launch_check("grid_nearest_backward_kernel");
		return;
	}
	GRID_DISPATCH(grid_backward_kernel, plan, g, stream, gp, B, x, dL_denc, grid_grads, replicas, chunk_begin);
	launch_check("grid_backward_kernel");
}
// Adds the replica sums of the replicated levels among the encoding chunks [chunk_begin, chunk_end) to the gradient buffer.
static void backward_finish(const tcnn_b200_plan* plan, cudaStream_t stream, const GridParams& gp, __half* grid_grads, __half* replicas,
	uint32_t chunk_begin = 0, uint32_t chunk_end = 0xFFFFFFFFu) {
	if (plan->info.n_features_per_level == 1) { // cast of the fp32 copy, added to what the gradient buffer holds (zero after an Overwrite prepare)
		if (chunk_begin != 0) return;
		const uint32_t n = (uint32_t)plan->info.n_params_grid;
		grid_cast_gradient_kernel<<<(n + 255) / 256, 256, 0, stream>>>(n, (const float*)replicas, grid_grads, 1);
		launch_check("grid_cast_gradient_kernel");
		return;
	}
	if (!gp.rep_entries) return;
	const uint32_t F = plan->info.n_features_per_level, lpc = 8 / F;
	const uint32_t level_begin = std::min(chunk_begin * lpc, gp.n_levels);
	const uint32_t level_end = chunk_end > gp.n_levels ? gp.n_levels : std::min(chunk_end * lpc, gp.n_levels);
	uint32_t entry_begin = 0xFFFFFFFFu, entry_end = 0;
	for (uint32_t l = level_begin; l < level_end; ++l) {
		if (!gp.level[l].replicated) continue;
		entry_begin = std::min(entry_begin, gp.level[l].rep_offset);
		entry_end = std::max(entry_end, gp.level[l].rep_offset + gp.level[l].size);
	}
	if (entry_begin >= entry_end) return;
	ProfileScope prof(plan, stream, PK_GRID_BWD);
	const uint32_t pair_begin = entry_begin * F / 2, pair_end = entry_end * F / 2;
	grid_replica_reduce_kernel<<<(pair_end - pair_begin + 63) / 64, 64, 0, stream>>>(gp, F, replicas, grid_grads, pair_begin, pair_end);
	launch_check("grid_replica_reduce_kernel");
}
static void encode_backward(const tcnn_b200_plan* plan, cudaStream_t stream, uint32_t B, const float* x, const __half* dL_denc, __half* grid_grads, int zero_first, __half* replicas) {
	const GridParams gp = backward_params(plan, B, replicas != nullptr);
	backward_prepare(plan, stream, gp, grid_grads, zero_first, replicas);
	backward_scatter(plan, stream, gp, B, x, dL_denc, grid_grads, replicas);
	backward_finish(plan, stream, gp, grid_grads, replicas);
}

static void encode_any(const tcnn_b200_plan* plan, cudaStream_t stream, uint32_t B, const float* x, const __half* grid, __half* encoded, bool image);
extern "C" int tcnn_b200_encode_forward(const tcnn_b200_plan* plan, void* stream, uint32_t B, const float* x, const void* grid_params, void* encoded) {
	API_BEGIN
	require_device(plan); check_batch(B);
	CHECK(x && encoded && (grid_params || plan->enc_kind != ENC_GRID), "null argument");
	CHECK(!plan->f32, "fp16 entry point: an fp32-precision encoding goes through tcnn_b200_forward / _inference / _backward");
	encode_any(plan, (cudaStream_t)stream, B, x, (const __half*)grid_params, (__half*)encoded, /*image=*/false); // OneBlob / Identity: grid_params unused
	API_END
}
extern "C" int tcnn_b200_encode_backward(const tcnn_b200_plan* plan, void* stream, uint32_t B, const float* x, const void* dL_denc, void* grid_grads, int zero_first, void* workspace) {
	API_BEGIN
	require_device(plan); check_batch(B);
	CHECK(x && dL_denc && grid_grads, "null argument");
	CHECK(!plan->f32, "fp16 entry point: an fp32-precision encoding goes through tcnn_b200_forward / _inference / _backward");
	encode_backward(plan, (cudaStream_t)stream, B, x, (const __half*)dL_denc, (__half*)grid_grads, zero_first, workspace ? carve(plan, B, workspace).replicas : nullptr);
	API_END
}
// ---- allocation-free variant of tcnn_b200_encode_backward ---------------------------------------------------------
// For callers that cannot allocate device memory (CUDA graph capture, tight serving loops, harnesses that forbid
// cudaMalloc): the scratch of the scatter (the private gradient copies of the coarse levels) is a file-scope device array
// instead of a slice of the caller's workspace.  Every GPU that loads the library holds these bytes until exit.
// In host code `g_noalloc_scratch` is only a stand-in for the symbol; its device address comes from cudaGetSymbolAddress,
// once per device with that device current.  The array is shared by all calls and keeps the last call's contents: the
// part a call uses is cleared on the call's stream first, and calls that overlap on the GPU race on it.
constexpr size_t NOALLOC_SCRATCH_BYTES = TCNN_B200_NOALLOC_SCRATCH_BYTES; // 2 MiB (include/tcnn_b200.h)
static __device__ __align__(256) unsigned char g_noalloc_scratch[NOALLOC_SCRATCH_BYTES];

static __half* noalloc_scratch_address() {
	constexpr int MAX_DEVICES = 64;
	static std::mutex mutex;
	static void* address[MAX_DEVICES] = {};
	int device = 0;
	CUDA_OK(cudaGetDevice(&device));
	CHECK(device >= 0 && device < MAX_DEVICES, "device ordinal out of range");
	std::lock_guard<std::mutex> lock(mutex);
	if (!address[device]) {
		void* p = nullptr;
		CUDA_OK(cudaGetSymbolAddress(&p, g_noalloc_scratch));
		CHECK(p != nullptr, "cudaGetSymbolAddress returned a null address");
		address[device] = p;
	}
	return (__half*)address[device];
}

extern "C" int tcnn_b200_encode_backward_noalloc(const tcnn_b200_plan* plan, void* stream, uint32_t B, const float* x, const void* dL_denc, void* grid_grads, int zero_first) {
	API_BEGIN
	CHECK(plan, "null plan");
	// require_device() would evaluate the level scales on first use and allocate for it; a plan created with a device present has them
	CHECK(plan->device_scales, "allocation-free entry point: the plan must have been created with a CUDA device present");
	CHECK(!plan->profiling, "allocation-free entry point: per-kernel profiling creates CUDA events, disable it");
	check_batch(B);
	CHECK(x && dL_denc && grid_grads, "null argument");
	CHECK(plan->enc_kind == ENC_GRID && !plan->f32, "allocation-free entry point: fp16 grid encodings only");
	const GridParams gp = backward_params(plan, B, true);
	const size_t need = plan->info.n_features_per_level == 1 ? (size_t)plan->info.n_params_grid * 4 : replica_bytes_of(plan, gp);
	CHECK(need <= NOALLOC_SCRATCH_BYTES, "allocation-free entry point: this grid needs " + std::to_string(need) + " bytes of scatter scratch, the built-in array holds " +
		std::to_string(NOALLOC_SCRATCH_BYTES) + " (use tcnn_b200_encode_backward with a workspace)");
	__half* scratch = noalloc_scratch_address();
	// backward_prepare clears the `need` bytes of the array this call accumulates into (cudaMemsetAsync on `stream`)
	encode_backward(plan, (cudaStream_t)stream, B, x, (const __half*)dL_denc, (__half*)grid_grads, zero_first, scratch);
	API_END
}

extern "C" int tcnn_b200_grid_indices(const tcnn_b200_plan* plan, void* stream, uint32_t B, const float* x, uint32_t* indices) {
	API_BEGIN
	require_device(plan);
	CHECK(x && indices && B > 0, "bad argument");
	const dim3 g((B + 255) / 256, plan->info.n_levels, 1);
	const uint32_t D = plan->info.n_input_dims;
	if (D == 2) grid_indices_kernel<2><<<g, 256, 0, (cudaStream_t)stream>>>(plan->grid, B, x, indices);
	else if (D == 3) grid_indices_kernel<3><<<g, 256, 0, (cudaStream_t)stream>>>(plan->grid, B, x, indices);
	else grid_indices_kernel<4><<<g, 256, 0, (cudaStream_t)stream>>>(plan->grid, B, x, indices);
	launch_check("grid_indices_kernel");
	API_END
}

// ---- MLP ----------------------------------------------------------------------------------------------
#ifndef TCNN_B200_SCATTER_WARPGROUPS
#define TCNN_B200_SCATTER_WARPGROUPS 2
#endif
// cudaFuncAttributeMaxDynamicSharedMemorySize of a kernel, only ever raised, one record per kernel for the whole process:
// the attribute belongs to the function, not to the calling thread - a thread-local cache let torch's autograd thread lower
// the value another thread had set for a wider input (the same instantiation serves several input widths).
template <typename K>
static uint32_t ensure_dynamic_smem(K kernel, uint32_t smem) {
	static std::mutex mutex;
	static std::map<std::pair<int, const void*>, uint32_t> configured; // per device: one process may drive several GPUs
	int device = 0;
	CUDA_OK(cudaGetDevice(&device));
	std::lock_guard<std::mutex> lock(mutex);
	uint32_t& have = configured[std::make_pair(device, (const void*)kernel)];
	if (have < smem) {
		CUDA_OK(cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
		have = smem;
	}
	return have;
}
template <int MODE, int H, int FD = 0, int FF = 2, int NSG = 0, bool GEN = false, bool WIN = false>
static void launch_mlp_k(const tcnn_b200_plan* plan, cudaStream_t stream, const MlpArgs& a, uint32_t grid, const GridParams& gp, uint32_t smem) {
	const uint32_t configured = ensure_dynamic_smem(mlp_kernel<MODE, H, FD, FF, NSG, GEN, WIN>, smem);
	ProfileScope prof(plan, stream, PK_MLP);
	mlp_kernel<MODE, H, FD, FF, NSG, GEN, WIN><<<grid, 128 + 128 * NSG, smem, stream>>>(a, gp);
	{
		const cudaError_t e = cudaGetLastError();
		if (e != cudaSuccess) {
			cudaFuncAttributes fa{};
			const cudaError_t qe = cudaFuncGetAttributes(&fa, mlp_kernel<MODE, H, FD, FF, NSG, GEN, WIN>);
			int dev = -1; cudaGetDevice(&dev);
			FAIL(std::string("mlp_kernel launch failed: ") + cudaGetErrorString(e) + " (mode " + std::to_string(MODE) + ", hidden " + std::to_string(H) +
				", grid " + std::to_string(grid) + ", threads " + std::to_string(128 + 128 * NSG) + ", smem " + std::to_string(smem) + ", configured " + std::to_string(configured) +
				", attr query " + cudaGetErrorString(qe) + ", maxDynamicSharedSizeBytes " + std::to_string(fa.maxDynamicSharedSizeBytes) + ", static smem " + std::to_string(fa.sharedSizeBytes) +
				", regs " + std::to_string(fa.numRegs) + ", device " + std::to_string(dev) + ", stream " + std::to_string((uintptr_t)stream) + ")");
		}
	}
}
template <int MODE, int H, int FD = 0, int FF = 2>
static void launch_mlp_h(const tcnn_b200_plan* plan, cudaStream_t stream, const MlpArgs& a, uint32_t grid, const GridParams& gp) {
	constexpr bool train = MODE != MODE_INFER;
	const uint32_t smem = mlp_smem_bytes(H, (int)a.in_width, train);
	if constexpr (FD > 0) {
		// warp-specialised scatter (second warpgroup + hand-over slot) unless the slot would cost a resident CTA per SM
		const uint32_t per_sm = 233472u, smem_ws = smem + mlp_ws_slot_bytes((int)a.in_width);
		if (per_sm / (smem_ws + 1024u) >= std::min(2u, per_sm / (smem + 1024u)) && per_sm / (smem_ws + 1024u) >= 1u && !plan->no_warp_specialization) {
			launch_mlp_k<MODE, H, FD, FF, TCNN_B200_SCATTER_WARPGROUPS>(plan, stream, a, grid, gp, smem_ws);
			return;
		}
	}
	if constexpr (FD == 0) {
		// hidden activation other than ReLU: the kernel variant that reads it at run time (never the fused-scatter kernels)
		// inputs wider than the network (64 < E <= 128): the variant with a two-sub-tile input (activation read at run time)
		if (a.in_width > 64) { launch_mlp_k<MODE, H, 0, 2, 0, true, true>(plan, stream, a, grid, gp, smem); return; }
		if (a.activation != ACT_RELU || a.n_neurons < 64) { launch_mlp_k<MODE, H, 0, 2, 0, true>(plan, stream, a, grid, gp, smem); return; }
	}
	launch_mlp_k<MODE, H, FD, FF, 0>(plan, stream, a, grid, gp, smem);
}
// ---- 128 neurons: forward (+ loss) -> dgrad chain -> split-K weight gradients (mlp_wide_kernels.cuh) ----
template <typename K>
static void wide_set_smem(K kernel, uint32_t smem) { ensure_dynamic_smem(kernel, smem); }
template <int MODE>
static void launch_mlp_wide(const tcnn_b200_plan* plan, cudaStream_t stream, const MlpArgs& a, uint32_t n_splits) {
	constexpr bool train = MODE != MODE_INFER;
	const uint32_t H = a.n_hidden, E = a.in_width;
	WideArgs w{};
	w.n_tiles = a.n_tiles; w.in_width = E; w.n_hidden = H; w.n_out = a.n_out; w.batch = (uint64_t)a.n_tiles * MLP_TILE;
	w.input = a.input; w.params = a.params; w.output = a.output; w.output_f32 = a.output_f32; w.target = a.target;
	w.hidden = a.wide_hidden; w.gtmp = a.wide_gtmp; w.dL_dinput = a.dL_dinput;
	w.dL_doutput = MODE == MODE_TRAIN ? a.wide_dy : a.dL_doutput; w.dL_doutput_out = a.wide_dy;
	w.activation = a.activation; w.output_activation = a.output_activation;
	w.dW_partials = a.dW_partials; w.loss_partials = a.loss_partials; w.n_mlp_params = a.n_mlp_params; w.n_splits = n_splits;
	w.loss_type = a.loss_type; w.loss_scale = a.loss_scale; w.n_total = a.n_total;
	if (train) CHECK(w.hidden && w.gtmp && w.dL_doutput, "the 128-neuron training / backward pass needs the workspace");
	{
		ProfileScope prof(plan, stream, PK_MLP);
		const uint32_t nwg = wide_nwg_fwd(plan), grid = wide_fwd_ctas(plan, a.n_tiles), smem = wide_fwd_smem(H, E, nwg);
		if (nwg == 2) { wide_set_smem(mlp_wide_forward_kernel<MODE, 2>, smem); mlp_wide_forward_kernel<MODE, 2><<<grid, 256, smem, stream>>>(w); }
		else { wide_set_smem(mlp_wide_forward_kernel<MODE, 1>, smem); mlp_wide_forward_kernel<MODE, 1><<<grid, 128, smem, stream>>>(w); }
		launch_check("mlp_wide_forward_kernel");
	}
	if constexpr (train) {
		if (MODE == MODE_BACKWARD && a.output_activation != ACT_NONE) w.dL_doutput = a.wide_dy; // written by the forward recomputation
		const bool want_dx = a.dL_dinput != nullptr;
		{
			ProfileScope prof(plan, stream, PK_MLP);
			const uint32_t nwg = wide_nwg_bwd(plan, want_dx), smem = wide_bwd_smem(H, E, nwg, want_dx);
			const uint32_t grid = std::max(1u, std::min((a.n_tiles + nwg - 1) / nwg, (uint32_t)plan->n_sms));
			if (nwg == 2) { wide_set_smem(mlp_wide_backward_kernel<2>, smem); mlp_wide_backward_kernel<2><<<grid, 256, smem, stream>>>(w); }
			else { wide_set_smem(mlp_wide_backward_kernel<1>, smem); mlp_wide_backward_kernel<1><<<grid, 128, smem, stream>>>(w); }
			launch_check("mlp_wide_backward_kernel");
		}
		if (a.dW_partials) {
			ProfileScope prof(plan, stream, PK_MLP);
			CHECK(n_splits >= 1 && n_splits <= a.n_tiles, "bad weight-gradient split count");
			wide_set_smem(mlp_wide_dw_kernel, WIDE_DW_SMEM);
			mlp_wide_dw_kernel<<<dim3(n_splits, H + 1, 1), 128, WIDE_DW_SMEM, stream>>>(w);
			launch_check("mlp_wide_dw_kernel");
		}
	}
}

template <int MODE>
static void launch_mlp(const tcnn_b200_plan* plan, cudaStream_t stream, const MlpArgs& a, uint32_t grid) {
	if (is_wide(plan)) { launch_mlp_wide<MODE>(plan, stream, a, grid); return; }
	switch (a.n_hidden) {
		case 1: launch_mlp_h<MODE, 1>(plan, stream, a, grid, plan->grid); break;
		case 2: launch_mlp_h<MODE, 2>(plan, stream, a, grid, plan->grid); break;
		case 3: launch_mlp_h<MODE, 3>(plan, stream, a, grid, plan->grid); break;
		case 4: launch_mlp_h<MODE, 4>(plan, stream, a, grid, plan->grid); break;
		case 5: launch_mlp_h<MODE, 5>(plan, stream, a, grid, plan->grid); break;
		default: FAIL("n_hidden_layers out of range");
	}
}
// Training kernel with the hash-table gradient scatter fused into its last stage; instantiated for the common shapes.
static bool fused_scatter_available(const tcnn_b200_plan* plan) {
	const uint32_t D = plan->info.n_input_dims, F = plan->info.n_features_per_level;
	return plan->info.n_neurons == 64 && plan->enc_kind == ENC_GRID && !general_grid(plan) && !plan->grid.smoothstep && !plan->grid.nearest && !plan->grid.stochastic && plan->activation == ACT_RELU && plan->info.n_hidden_layers <= 4
		&& plan->info.encoded_width <= 64 && (D == 2 || D == 3) && (F == 2 || F == 4);
}
template <int H>
static void launch_mlp_train_fused_h(const tcnn_b200_plan* plan, cudaStream_t stream, const MlpArgs& a, uint32_t grid, const GridParams& gp) {
	const uint32_t D = plan->info.n_input_dims, F = plan->info.n_features_per_level;
	if (D == 2 && F == 2) launch_mlp_h<MODE_TRAIN, H, 2, 2>(plan, stream, a, grid, gp);
	else if (D == 3 && F == 2) launch_mlp_h<MODE_TRAIN, H, 3, 2>(plan, stream, a, grid, gp);
	else if (D == 2 && F == 4) launch_mlp_h<MODE_TRAIN, H, 2, 4>(plan, stream, a, grid, gp);
	else if (D == 3 && F == 4) launch_mlp_h<MODE_TRAIN, H, 3, 4>(plan, stream, a, grid, gp);
	else FAIL("no fused-scatter kernel for this (n_input_dims, n_features_per_level)");
}
static void launch_mlp_train_fused(const tcnn_b200_plan* plan, cudaStream_t stream, const MlpArgs& a, uint32_t grid, const GridParams& gp) {
	switch (a.n_hidden) {
		case 1: launch_mlp_train_fused_h<1>(plan, stream, a, grid, gp); break;
		case 2: launch_mlp_train_fused_h<2>(plan, stream, a, grid, gp); break;
		case 3: launch_mlp_train_fused_h<3>(plan, stream, a, grid, gp); break;
		case 4: launch_mlp_train_fused_h<4>(plan, stream, a, grid, gp); break;
		default: FAIL("n_hidden_layers out of range");
	}
}

static MlpArgs base_mlp_args(const tcnn_b200_plan* plan, uint32_t B, const __half* input, const __half* params, const Workspace* w = nullptr) {
	MlpArgs a{};
	if (w) { a.wide_hidden = w->wide_hidden; a.wide_gtmp = w->wide_gtmp; a.wide_dy = w->wide_dy; }
	a.n_tiles = B / MLP_TILE;
	a.in_width = plan->info.encoded_width;
	a.n_hidden = plan->info.n_hidden_layers;
	a.n_out = plan->info.n_output_dims;
	a.n_neurons = plan->info.n_neurons;
	a.input = input;
	a.params = params;
	a.n_mlp_params = (uint32_t)plan->info.n_params_mlp;
	a.loss_type = plan->info.loss_type;
	a.n_total = (float)((uint64_t)B * plan->info.n_output_dims);
	a.loss_scale = 128.f;
	a.activation = plan->activation; a.output_activation = plan->output_activation;
	return a;
}

static void mlp_infer(const tcnn_b200_plan* plan, cudaStream_t stream, uint32_t B, const __half* input, const __half* params, __half* out_h, float* out_f) {
	MlpArgs a = base_mlp_args(plan, B, input, params);
	a.output = out_h;
	a.output_f32 = out_f;
	launch_mlp<MODE_INFER>(plan, stream, a, mlp_grid_size(plan, a.n_tiles, false));
}

static void mlp_finalize(const tcnn_b200_plan* plan, cudaStream_t stream, uint32_t n_parts, const Workspace& w, __half* grads, float* loss_out, bool with_loss, uint32_t n_tiles = 0,
	bool accumulate = false) {
	ProfileScope prof(plan, stream, PK_FINALIZE);
	const uint32_t n = (uint32_t)plan->info.n_params_mlp;
	mlp_finalize_kernel<<<(n + 31) / 32, 256, 0, stream>>>(n, n_parts, w.dw_partials, grads, with_loss ? w.loss_partials : nullptr, loss_out,
		mlp_loss_parts(plan, n_tiles, n_parts), accumulate ? 1 : 0);
	launch_check("mlp_finalize_kernel");
}

static void rows_to_image(const tcnn_b200_plan* plan, cudaStream_t stream, uint32_t B, const __half* rows, __half* image) {
	const uint32_t n = B * (plan->info.encoded_width / 8);
	wide_rowmajor_to_image_kernel<<<(n + 255) / 256, 256, 0, stream>>>(B, plan->info.encoded_width, rows, image);
	launch_check("wide_rowmajor_to_image_kernel");
}

extern "C" int tcnn_b200_mlp_forward(const tcnn_b200_plan* plan, void* stream, uint32_t B, const void* input, const void* params, void* output) {
	API_BEGIN
	require_device(plan); check_batch(B);
	CHECK(input && params && output, "null argument");
	if (is_wide(plan)) { // the 128-neuron kernels read tile images: convert the caller's rows into a stream-ordered scratch buffer
		cudaStream_t st = (cudaStream_t)stream;
		void* img = nullptr;
		CUDA_OK(cudaMallocAsync(&img, encoded_bytes(plan, B), st));
		rows_to_image(plan, st, B, (const __half*)input, (__half*)img);
		mlp_infer(plan, st, B, (const __half*)img, (const __half*)params, (__half*)output, nullptr);
		CUDA_OK(cudaFreeAsync(img, st));
		return 0;
	}
	mlp_infer(plan, (cudaStream_t)stream, B, (const __half*)input, (const __half*)params, (__half*)output, nullptr);
	API_END
}

extern "C" int tcnn_b200_mlp_backward(const tcnn_b200_plan* plan, void* stream_, uint32_t B, const void* input, const void* params,
	const void* dL_doutput, void* dL_dparams, void* dL_dinput, void* workspace) {
	API_BEGIN
	require_device(plan); check_batch(B);
	CHECK(input && params && dL_doutput, "null argument");
	CHECK(!dL_dparams || workspace, "workspace required for weight gradients");
	cudaStream_t stream = (cudaStream_t)stream_;
	Workspace w{};
	if (workspace) w = carve(plan, B, workspace);
	MlpArgs a = base_mlp_args(plan, B, (const __half*)input, (const __half*)params, workspace ? &w : nullptr);
	if (is_wide(plan)) {
		CHECK(workspace, "the 128-neuron backward pass needs the workspace");
		rows_to_image(plan, stream, B, (const __half*)input, w.encoded);
		a.input = w.encoded;
	}
	a.dL_doutput = (const __half*)dL_doutput;
	a.dL_dinput = (__half*)dL_dinput;
	a.dW_partials = dL_dparams ? w.dw_partials : nullptr;
	const uint32_t grid = mlp_grid_size(plan, a.n_tiles, true);
	launch_mlp<MODE_BACKWARD>(plan, stream, a, grid);
	if (dL_dparams) mlp_finalize(plan, stream, grid, w, (__half*)dL_dparams, nullptr, false);
	API_END
}

// ---- encoding + MLP -------------------------------------------------------------------------------------
static void encode_backward_input(const tcnn_b200_plan* plan, cudaStream_t stream, uint32_t B, const float* x, const __half* grid, const __half* dL_denc, float* dL_dx) {
	ProfileScope prof(plan, stream, PK_GRID_BWD);
	const dim3 g((B + 255) / 256, 1, 1);
	if (plan->grid.nearest) { // "Gradient is zero when there's no interpolation" (grid.h:131-137)
		CUDA_OK(cudaMemsetAsync(dL_dx, 0, (size_t)B * plan->info.n_input_dims * sizeof(float), stream));
		return;
	}
	GRID_DISPATCH_F1(grid_backward_input_kernel, plan, g, stream, plan->grid, B, x, grid, dL_denc, dL_dx);
	launch_check("grid_backward_input_kernel");
}

// grid gather, or the Identity encoding of a stand-alone Network
// `image`: write tile images for the 128-neuron MLP instead of rows (internal callers pass is_wide(plan))
static void encode_any(const tcnn_b200_plan* plan, cudaStream_t stream, uint32_t B, const float* x, const __half* grid, __half* encoded, bool image) {
	if (plan->enc_kind == ENC_GRID) { encode_forward(plan, stream, B, x, grid, encoded, nullptr, 0, image); return; }
	const uint32_t subs = image ? image_subs(plan) : 0u;
	ProfileScope prof(plan, stream, PK_GRID_FWD);
	if (plan->enc_kind == ENC_ONEBLOB) {
		const uint32_t n_out = plan->grid.n_features, rows = (128 + n_out - 1) / n_out; // the reference's block shape (oneblob.h:118-121)
		const uint32_t per_block = rows * ONEBLOB_ROUNDS;
		oneblob_encode_kernel<<<(B + per_block - 1) / per_block, dim3(n_out, rows, 1), 0, stream>>>(B, plan->info.n_input_dims, plan->oneblob_log2_bins, n_out,
			plan->info.encoded_width, x, encoded, subs);
		launch_check("oneblob_encode_kernel");
		return;
	}
	const uint32_t n = B * (plan->info.encoded_width / 8);
	identity_encode_kernel<<<(n + 255) / 256, 256, 0, stream>>>(B, plan->info.n_input_dims, plan->info.encoded_width, x, encoded, subs);
	launch_check("identity_encode_kernel");
}

extern "C" int tcnn_b200_inference(const tcnn_b200_plan* plan, void* stream_, uint32_t B, const float* x, const void* params,
	void* output_fp16, float* output_fp32, void* workspace) {
	API_BEGIN
	require_device(plan); check_batch(B);
	CHECK(x && params && workspace && (output_fp16 || output_fp32), "null argument");
	cudaStream_t stream = (cudaStream_t)stream_;
	const Workspace w = carve(plan, B, workspace);
	const __half* p = (const __half*)params;
	if (plan->kind == PLAN_ENCODING && plan->f32) { // fp32 table -> fp32 [B][encoded_width] (either output pointer)
		CHECK(plan->enc_kind == ENC_GRID, "fp32 precision is implemented for grid encodings");
		float* out = output_fp32 ? output_fp32 : (float*)output_fp16;
		ProfileScope prof(plan, stream, PK_GRID_FWD);
		const dim3 g((B + 255) / 256, std::max(plan->grid.n_levels, 1u), 1);
		GRID_DISPATCH(grid_forward_f32_kernel, plan, g, stream, plan->grid, B, x, (const float*)params, out);
		launch_check("grid_forward_f32_kernel");
		return 0;
	}
	if (plan->kind == PLAN_ENCODING) { // tcnn.Encoding: output = the encoded batch, fp16 [B][encoded_width]
		CHECK(output_fp16 && !output_fp32, "a stand-alone encoding writes fp16 [B][encoded_width]");
		encode_forward(plan, stream, B, x, p, (__half*)output_fp16);
		return 0;
	}
	encode_any(plan, stream, B, x, p + plan->info.n_params_mlp, w.encoded, is_wide(plan));
	mlp_infer(plan, stream, B, w.encoded, p, (__half*)output_fp16, output_fp32);
	API_END
}

extern "C" int tcnn_b200_forward(const tcnn_b200_plan* plan, void* stream, uint32_t B, const float* x, const void* params, void* output_fp16, void* workspace) {
	return tcnn_b200_inference(plan, stream, B, x, params, output_fp16, nullptr, workspace);
}

// GradientMode (object.h:115-119)
enum GradientModeKind { GRADIENT_IGNORE = 0, GRADIENT_OVERWRITE = 1, GRADIENT_ACCUMULATE = 2 };
extern "C" int tcnn_b200_backward_mode(const tcnn_b200_plan* plan, void* stream_, uint32_t B, const float* x, const void* params,
	const void* dL_doutput, void* dL_dparams, float* dL_dinput, void* workspace, int gradient_mode) {
	API_BEGIN
	require_device(plan); check_batch(B);
	CHECK(x && params && dL_doutput && workspace, "null argument");
	CHECK(gradient_mode >= GRADIENT_IGNORE && gradient_mode <= GRADIENT_ACCUMULATE, "gradient_mode must be 0 (Ignore), 1 (Overwrite) or 2 (Accumulate)");
	if (gradient_mode == GRADIENT_IGNORE) dL_dparams = nullptr;
	const bool accumulate = gradient_mode == GRADIENT_ACCUMULATE;
	if (!dL_dparams && !dL_dinput) return 0;
	cudaStream_t stream = (cudaStream_t)stream_;
	const Workspace w = carve(plan, B, workspace);
	const __half* p = (const __half*)params;
	__half* g = (__half*)dL_dparams;
	if (plan->kind == PLAN_ENCODING && plan->f32) { // everything fp32: dL/d(encoded) [B][encoded_width], table gradient (GradientMode::Overwrite)
		ProfileScope prof(plan, stream, PK_GRID_BWD);
		if (dL_dparams) {
			if (!accumulate) CUDA_OK(cudaMemsetAsync(dL_dparams, 0, plan->info.n_params_grid * sizeof(float), stream)); // grid.h:857-859
			const dim3 gl((B + 255) / 256, std::max(plan->grid.n_levels, 1u), 1);
			GRID_DISPATCH(grid_backward_f32_kernel, plan, gl, stream, plan->grid, B, x, (const float*)dL_doutput, (float*)dL_dparams);
			launch_check("grid_backward_f32_kernel");
		}
		if (dL_dinput) {
			if (plan->grid.nearest) CUDA_OK(cudaMemsetAsync(dL_dinput, 0, (size_t)B * plan->info.n_input_dims * sizeof(float), stream));
			else {
				const dim3 gs((B + 255) / 256, 1, 1);
				GRID_DISPATCH_T(grid_backward_input_kernel, float, plan, gs, stream, plan->grid, B, x, (const float*)params, (const float*)dL_doutput, dL_dinput);
				launch_check("grid_backward_input_kernel<float>");
			}
		}
		return 0;
	}
	if (plan->kind == PLAN_ENCODING) { // dL_doutput is dL/d(encoded), fp16 [B][encoded_width]
		if (g) encode_backward(plan, stream, B, x, (const __half*)dL_doutput, g, /*zero_first=*/accumulate ? 0 : 1, w.replicas);
		if (dL_dinput) encode_backward_input(plan, stream, B, x, p, (const __half*)dL_doutput, dL_dinput);
		return 0;
	}
	MlpArgs a = base_mlp_args(plan, B, w.encoded, p, &w); // w.encoded was left by tcnn_b200_forward (the Context)
	a.dL_doutput = (const __half*)dL_doutput;
	a.dL_dinput = w.dL_denc;
	a.dW_partials = g ? w.dw_partials : nullptr;
	const uint32_t grid = mlp_grid_size(plan, a.n_tiles, true);
	launch_mlp<MODE_BACKWARD>(plan, stream, a, grid);
	if (g) mlp_finalize(plan, stream, grid, w, g, nullptr, false, 0, accumulate); // beta = 1 of the weight-gradient GEMMs (fully_fused_mlp.cu:769)
	if (plan->enc_kind == ENC_IDENTITY) {
		if (dL_dinput) {
			const uint32_t n = B * plan->info.n_input_dims;
			identity_backward_kernel<<<(n + 255) / 256, 256, 0, stream>>>(B, plan->info.n_input_dims, plan->info.encoded_width, w.dL_denc, dL_dinput);
			launch_check("identity_backward_kernel");
		}
	} else if (plan->enc_kind == ENC_ONEBLOB) {
		if (dL_dinput) {
			const uint32_t n = B * plan->info.n_input_dims;
			oneblob_backward_input_kernel<<<(n + 255) / 256, 256, 0, stream>>>(B, plan->info.n_input_dims, plan->oneblob_log2_bins, plan->info.encoded_width, x, w.dL_denc, dL_dinput);
			launch_check("oneblob_backward_input_kernel");
		}
	} else {
		if (g) encode_backward(plan, stream, B, x, w.dL_denc, g + plan->info.n_params_mlp, /*zero_first=*/accumulate ? 0 : 1, w.replicas);
		if (dL_dinput) encode_backward_input(plan, stream, B, x, p + plan->info.n_params_mlp, w.dL_denc, dL_dinput);
	}
	API_END
}
extern "C" int tcnn_b200_backward(const tcnn_b200_plan* plan, void* stream, uint32_t B, const float* x, const void* params,
	const void* dL_doutput, void* dL_dparams, float* dL_dinput, void* workspace) {
	return tcnn_b200_backward_mode(plan, stream, B, x, params, dL_doutput, dL_dparams, dL_dinput, workspace, GRADIENT_OVERWRITE);
}

// DifferentiableObject::backward_backward_input (object.h:208-240) of the grid encoding (grid.h:902-1026).
extern "C" int tcnn_b200_backward_backward_input(const tcnn_b200_plan* plan, void* stream_, uint32_t B, const float* x, const void* params,
	const float* dL_ddLdinput, const void* dL_doutput, void* dL_ddLdoutput, void* dL_dparams, float* dL_dinput, void* workspace) {
	API_BEGIN
	require_device(plan); check_batch(B);
	CHECK(!plan->f32, "second-order gradients are implemented for fp16 precision");
	CHECK(plan->info.n_features_per_level != 1 && !plan->grid.max_level_gpu, "second-order gradients are implemented for n_features_per_level 2, 4, 8 without a per-sample max level");
	CHECK(plan->kind == PLAN_ENCODING && plan->enc_kind == ENC_GRID,
		"second-order gradients exist for the stand-alone grid encoding only (FullyFusedMLP has no backward_backward_input in the reference either)");
	CHECK(x && params && dL_ddLdinput && dL_doutput, "null argument");
	(void)workspace;
	cudaStream_t stream = (cudaStream_t)stream_;
	const dim3 g((B + 255) / 256, 1, 1);
	if (plan->grid.nearest) { // dy/dx == 0 everywhere (grid.h:420-423, 522-525 return before touching anything)
		if (dL_ddLdoutput) CUDA_OK(cudaMemsetAsync(dL_ddLdoutput, 0, (size_t)B * plan->info.encoded_width * 2, stream));
		if (dL_dinput) CUDA_OK(cudaMemsetAsync(dL_dinput, 0, (size_t)B * plan->info.n_input_dims * sizeof(float), stream));
		if (dL_dparams) CUDA_OK(cudaMemsetAsync(dL_dparams, 0, plan->info.n_params_grid * 2, stream));
		return 0;
	}
	if (dL_ddLdoutput || dL_dinput) {
		ProfileScope prof(plan, stream, PK_GRID_BWD);
		GRID_DISPATCH(grid_backward_backward_input_kernel, plan, g, stream, plan->grid, B, x, (const __half*)params, (const __half*)dL_doutput, dL_ddLdinput,
			(__half*)dL_ddLdoutput, dL_dinput);
		launch_check("grid_backward_backward_input_kernel");
	}
	if (dL_dparams) { // GradientMode::Overwrite
		ProfileScope prof(plan, stream, PK_GRID_BWD);
		CUDA_OK(cudaMemsetAsync(dL_dparams, 0, plan->info.n_params_grid * 2, stream));
		const dim3 gl((B + 255) / 256, std::max(plan->grid.n_levels, 1u), 1);
		GRID_DISPATCH(grid_backward_backward_grid_kernel, plan, gl, stream, plan->grid, B, x, (const __half*)dL_doutput, dL_ddLdinput, (__half*)dL_dparams);
		launch_check("grid_backward_backward_grid_kernel");
	}
	API_END
}

extern "C" int tcnn_b200_training_fwd_bwd(const tcnn_b200_plan* plan, void* stream_, uint32_t B, const float* x, const float* target,
	const void* params, void* grads, float loss_scale, int zero_grid_grads, float* loss_out, void* output_fp16, void* workspace) {
	API_BEGIN
	require_device(plan); check_batch(B);
	CHECK(plan->kind == PLAN_FULL, "training entry points need an encoding and a network");
	CHECK(x && target && params && grads && workspace, "null argument");
	cudaStream_t main = (cudaStream_t)stream_;
	const Workspace w = carve(plan, B, workspace);
	const tcnn_b200_info& in = plan->info;
	const __half* p = (const __half*)params;
	__half* g = (__half*)grads;
	__half* gg = g + in.n_params_mlp;
	const GridParams gp = backward_params(plan, B, true);

	// One stream, back to back.  A two-chunk pipeline over a high-priority stream (MLP of chunk k overlapping the gather of
	// chunk k+1 / scatter of chunk k-1) was measured and dropped (r01): the MLP CTAs need the max-shared carve-out and
	// cannot become resident next to gather blocks running with the max-L1 carve-out, and giving the gather kernel the
	// shared carve-out doubles its time.
	const uint32_t n_chunks = 1;
	uint32_t part_off[1] = {0}, part_cnt[1] = {mlp_grid_size(plan, B / MLP_TILE, true)};
	MlpArgs a = base_mlp_args(plan, B, w.encoded, p, &w);
	a.target = target;
	a.output = (__half*)output_fp16;
	a.dL_dinput = w.dL_denc;
	a.dW_partials = w.dw_partials;
	a.loss_partials = w.loss_partials;
	a.loss_scale = loss_scale;
	if (plan->enc_kind != ENC_GRID) { // parameter-free encoding (OneBlob): no table gradients, dL/d(encoded) is not needed
		encode_any(plan, main, B, x, nullptr, w.encoded, is_wide(plan));
		a.dL_dinput = nullptr;
		launch_mlp<MODE_TRAIN>(plan, main, a, part_cnt[0]);
		mlp_finalize(plan, main, part_cnt[0], w, g, loss_out, loss_out != nullptr, B / MLP_TILE);
		return 0;
	}
	backward_prepare(plan, main, gp, gg, zero_grid_grads, w.replicas);
	encode_forward(plan, main, B, x, p + in.n_params_mlp, w.encoded, nullptr, 0, is_wide(plan));
	if (fused_scatter_available(plan)) {
		a.dL_dinput = nullptr; a.positions = x; a.grid_grad = gg; a.replicas = w.replicas;
		launch_mlp_train_fused(plan, main, a, part_cnt[0], gp);
	} else {
		launch_mlp<MODE_TRAIN>(plan, main, a, part_cnt[0]);
		backward_scatter(plan, main, gp, B, x, w.dL_denc, gg, w.replicas);
	}
	mlp_finalize(plan, main, part_off[n_chunks - 1] + part_cnt[n_chunks - 1], w, g, loss_out, loss_out != nullptr, B / MLP_TILE);
	backward_finish(plan, main, gp, gg, w.replicas);
	API_END
}

// ---- optimizer --------------------------------------------------------------------------------------------
struct AdamBuffers { float* master; __half* params; __half* grads; float* m; float* v; uint32_t* steps; };
// the members of AdamArgs that change from one optimizer step to the next (everything else is fixed per plan and buffers)
static void set_adam_step(const tcnn_b200_plan* plan, AdamArgs& a, uint32_t adam_step_count) {
	const tcnn_b200_info& in = plan->info;
	a.lower_lr_bound = 0.f; a.upper_lr_bound = std::numeric_limits<float>::max();
	if (plan->adabound) { // optimizers/adam.h:156-160
		a.lower_lr_bound = 0.1f - 0.1f / ((1 - in.beta2) * (float)adam_step_count + 1);
		a.upper_lr_bound = 0.1f + 0.1f / ((1 - in.beta2) * (float)adam_step_count);
	}
	a.global_step = adam_step_count;
	a.global_correction = std::sqrt(1 - std::pow(in.beta2, (float)adam_step_count)) / (1 - std::pow(in.beta1, (float)adam_step_count));
}
static AdamArgs adam_args(const tcnn_b200_plan* plan, float grad_divisor, uint32_t adam_step_count, const AdamBuffers& b,
	int zero_grid_grads, uint32_t begin, uint32_t end) {
	const tcnn_b200_info& in = plan->info;
	AdamArgs a{};
	a.n = std::min<uint32_t>(end, (uint32_t)in.n_params);
	a.begin = begin;
	a.n_matrix = (uint32_t)in.n_params_mlp;
	a.loss_scale = grad_divisor;
	a.learning_rate = in.learning_rate; a.beta1 = in.beta1; a.beta2 = in.beta2; a.epsilon = in.epsilon; a.l2_reg = in.l2_reg;
	a.relative_decay = in.relative_decay; a.absolute_decay = in.absolute_decay;
	set_adam_step(plan, a, adam_step_count);
	a.weight_clip = plan->weight_clip;
	a.non_matrix_lr_factor = plan->non_matrix_lr_factor;
	a.optimize_matrix = plan->optimize_matrix_params ? 1 : 0; a.optimize_non_matrix = plan->optimize_non_matrix_params ? 1 : 0;
	a.zero_grads = zero_grid_grads;
	int e = 0;
	const float mant = std::frexp(grad_divisor, &e);
	a.inv_loss_scale = (mant == 0.5f) ? 1.0f / grad_divisor : 0.f; // exact only for powers of two
	a.log2_beta1 = std::log2(in.beta1); a.log2_beta2 = std::log2(in.beta2);
	a.master = b.master; a.params = b.params; a.grads = b.grads; a.m = b.m; a.v = b.v; a.steps = b.steps;
	return a;
}
// true if the replicated levels of gp are a prefix of the grid table laid out like the table itself (rep_offset == offset):
// then adam_kernel can sum the replicas while it reads the gradients (AdamArgs::replicas)
static bool replicas_are_table_prefix(const GridParams& gp) {
	bool ended = false;
	for (uint32_t l = 0; l < gp.n_levels; ++l) {
		const GridLevel& lv = gp.level[l];
		if (!lv.replicated) { ended = true; continue; }
		if (ended || lv.rep_offset != lv.offset) return false;
	}
	return true;
}
static void adam_launch(const tcnn_b200_plan* plan, cudaStream_t stream, float grad_divisor, uint32_t adam_step_count, const AdamBuffers& b,
	int zero_grid_grads, uint32_t begin, uint32_t end, const GridParams* reduce_replicas_of = nullptr, const __half* replicas = nullptr) {
	AdamArgs a = adam_args(plan, grad_divisor, adam_step_count, b, zero_grid_grads, begin, end);
	if (reduce_replicas_of && reduce_replicas_of->rep_entries) {
		a.replicas = replicas; a.n_replicas = reduce_replicas_of->n_replicas;
		a.rep_begin = (uint32_t)plan->info.n_params_mlp;
		a.rep_end = a.rep_begin + reduce_replicas_of->rep_entries * plan->info.n_features_per_level;
	}
	if (a.begin >= a.n) return;
	ProfileScope prof(plan, stream, PK_ADAM);
	const uint32_t n_threads = (a.n - a.begin + 3) / 4;
	adam_kernel<<<(n_threads + 255) / 256, 256, 0, stream>>>(a);
	launch_check("adam_kernel");
}

extern "C" int tcnn_b200_adam_step(const tcnn_b200_plan* plan, void* stream, float grad_divisor, uint32_t adam_step_count,
	float* params_fp32, void* params_fp16, void* grads_fp16, float* first_moments, float* second_moments, uint32_t* param_steps, int zero_grid_grads) {
	API_BEGIN
	require_device(plan);
	CHECK(params_fp32 && params_fp16 && grads_fp16 && first_moments && second_moments && param_steps, "null argument");
	adam_launch(plan, (cudaStream_t)stream, grad_divisor, adam_step_count,
		AdamBuffers{params_fp32, (__half*)params_fp16, (__half*)grads_fp16, first_moments, second_moments, param_steps}, zero_grid_grads, 0, (uint32_t)plan->info.n_params);
	API_END
}

extern "C" int tcnn_b200_adam_step_range(const tcnn_b200_plan* plan, void* stream, float grad_divisor, uint32_t adam_step_count,
	float* params_fp32, void* params_fp16, void* grads_fp16, float* first_moments, float* second_moments, uint32_t* param_steps, int zero_grid_grads,
	uint64_t param_begin, uint64_t param_end) {
	API_BEGIN
	require_device(plan);
	CHECK(params_fp32 && params_fp16 && grads_fp16 && first_moments && second_moments && param_steps, "null argument");
	CHECK(param_begin % 4 == 0 && param_begin <= param_end && param_end <= plan->info.n_params, "bad parameter range");
	adam_launch(plan, (cudaStream_t)stream, grad_divisor, adam_step_count,
		AdamBuffers{params_fp32, (__half*)params_fp16, (__half*)grads_fp16, first_moments, second_moments, param_steps}, zero_grid_grads, (uint32_t)param_begin, (uint32_t)param_end);
	API_END
}

// Fused Trainer::training_step (trainer.h:163-190) for one GPU: forward, loss, backward and the Adam update in one call.
// The scatter runs in level groups from the finest (largest tables) to the coarsest; as soon as a group's gradients are
// complete its Adam update starts on a second stream.  The scatter is bound by L2 atomic throughput (ncu r01b: lts 68 %,
// DRAM 5 %), Adam by HBM (DRAM 63 %), so the two overlap almost perfectly.  Requires zeroed grid gradients on entry
// (tcnn_b200_adam_step / this function leave them zeroed).
static void training_step_launches(const tcnn_b200_plan* plan, cudaStream_t main, uint32_t B, const float* x, const float* target,
	float loss_scale, uint32_t adam_step_count, float* params_fp32, void* params_fp16, void* grads_fp16, float* first_moments, float* second_moments,
	uint32_t* param_steps, float* loss_out, void* output_fp16, void* workspace) {
	cudaStream_t aux = plan->aux;
	const cudaEvent_t* ev = plan->ev;
	const Workspace w = carve(plan, B, workspace);
	const tcnn_b200_info& in = plan->info;
	const __half* p = (const __half*)params_fp16;
	__half* g = (__half*)grads_fp16;
	__half* gg = g + in.n_params_mlp;
	const AdamBuffers ab{params_fp32, (__half*)params_fp16, g, first_moments, second_moments, param_steps};
	const GridParams gp = backward_params(plan, B, true);
	const uint32_t F = in.n_features_per_level, levels_per_chunk = F ? 8 / F : 0;

	MlpArgs a = base_mlp_args(plan, B, w.encoded, p, &w);
	a.target = target; a.output = (__half*)output_fp16; a.dL_dinput = w.dL_denc;
	a.dW_partials = w.dw_partials; a.loss_partials = w.loss_partials; a.loss_scale = loss_scale;
	const uint32_t mlp_grid = mlp_grid_size(plan, B / MLP_TILE, true);

	if (fused_scatter_available(plan)) {
		// gather (+ clears the gradient replicas) -> [MLP forward, loss, backward, weight gradients, hash-table scatter] ->
		// replica reduction + Adam over the grid tables || on the second stream: weight-gradient reduction + Adam of the MLP weights
		encode_forward(plan, main, B, x, p + in.n_params_mlp, w.encoded, w.replicas, replica_bytes_of(plan, gp));
		a.dL_dinput = nullptr; a.positions = x; a.grid_grad = gg; a.replicas = w.replicas;
		launch_mlp_train_fused(plan, main, a, mlp_grid, gp);
		// while the per-kernel timer is on, the small kernel runs in line: on the second stream its event pair would measure how long
		// it waits for SMs next to the table-wide Adam kernel, not its own duration
		cudaStream_t side = plan->profiling ? main : aux;
		CUDA_OK(cudaEventRecord(ev[0], main));
		CUDA_OK(cudaStreamWaitEvent(side, ev[0], 0));
		{
			ProfileScope prof(plan, side, PK_FINALIZE);
			const AdamArgs am = adam_args(plan, loss_scale, adam_step_count, ab, 1, 0u, (uint32_t)in.n_params_mlp);
			mlp_finalize_adam_kernel<<<((uint32_t)in.n_params_mlp + 31) / 32, 256, 0, side>>>((uint32_t)in.n_params_mlp, mlp_grid, w.dw_partials,
				loss_out ? w.loss_partials : nullptr, loss_out, am, mlp_grid);
			launch_check("mlp_finalize_adam_kernel");
		}
		CUDA_OK(cudaEventRecord(ev[1], side));
		if (replicas_are_table_prefix(gp) && in.n_params_mlp % 4 == 0) { // replica sums inside the optimizer kernel
			adam_launch(plan, main, loss_scale, adam_step_count, ab, 1, (uint32_t)in.n_params_mlp, (uint32_t)in.n_params, &gp, w.replicas);
		} else {
			backward_finish(plan, main, gp, gg, w.replicas);
			adam_launch(plan, main, loss_scale, adam_step_count, ab, 1, (uint32_t)in.n_params_mlp, (uint32_t)in.n_params);
		}
		CUDA_OK(cudaStreamWaitEvent(main, ev[1], 0)); // join
		return;
	}
	if (plan->enc_kind != ENC_GRID) { // parameter-free encoding: encode -> MLP forward/backward -> weight-gradient reduction + Adam in one kernel
		encode_any(plan, main, B, x, nullptr, w.encoded, is_wide(plan));
		a.dL_dinput = nullptr;
		launch_mlp<MODE_TRAIN>(plan, main, a, mlp_grid);
		ProfileScope prof(plan, main, PK_FINALIZE);
		const AdamArgs am = adam_args(plan, loss_scale, adam_step_count, ab, 1, 0u, (uint32_t)in.n_params_mlp);
		mlp_finalize_adam_kernel<<<((uint32_t)in.n_params_mlp + 31) / 32, 256, 0, main>>>((uint32_t)in.n_params_mlp, mlp_grid, w.dw_partials,
			loss_out ? w.loss_partials : nullptr, loss_out, am, mlp_loss_parts(plan, B / MLP_TILE, mlp_grid));
		launch_check("mlp_finalize_adam_kernel");
		return;
	}
	backward_prepare(plan, main, gp, gg, /*zero_first=*/0, w.replicas);
	encode_forward(plan, main, B, x, p + in.n_params_mlp, w.encoded, nullptr, 0, is_wide(plan));
	launch_mlp<MODE_TRAIN>(plan, main, a, mlp_grid);
	CUDA_OK(cudaEventRecord(ev[0], main));
	CUDA_OK(cudaStreamWaitEvent(aux, ev[0], 0));
	mlp_finalize(plan, aux, mlp_grid, w, g, loss_out, loss_out != nullptr, B / MLP_TILE); // under the scatter

	// level groups: up to 4, finest first
	const uint32_t n_chunks = n_grad_chunks(plan);
	const uint32_t n_groups = general_grid(plan) ? 1u : std::min(4u, n_chunks); // the general scatter covers all levels in one launch
	uint32_t group_end = n_chunks;
	for (uint32_t gi = 0; gi < n_groups; ++gi) {
		const uint32_t remaining_groups = n_groups - gi;
		const uint32_t group_begin = group_end - (group_end + remaining_groups - 1) / remaining_groups;
		backward_scatter(plan, main, gp, B, x, w.dL_denc, gg, w.replicas, group_begin, group_end);
		const bool last = group_begin == 0;
		backward_finish(plan, main, gp, gg, w.replicas, group_begin, group_end); // replica sums of this group's coarse levels
		CUDA_OK(cudaEventRecord(ev[1 + gi], main));
		CUDA_OK(cudaStreamWaitEvent(aux, ev[1 + gi], 0));
		const uint32_t lvl_begin = std::min(group_begin * levels_per_chunk, in.n_levels), lvl_end = std::min(group_end * levels_per_chunk, in.n_levels);
		const uint32_t pb = (uint32_t)in.n_params_mlp + in.level_offset[lvl_begin] * F, pe = (uint32_t)in.n_params_mlp + in.level_offset[lvl_end] * F;
		adam_launch(plan, aux, loss_scale, adam_step_count, ab, 1, last ? 0u : pb, pe); // the last group also takes the MLP weights
		group_end = group_begin;
	}
	CUDA_OK(cudaEventRecord(ev[6], aux));
	CUDA_OK(cudaStreamWaitEvent(main, ev[6], 0)); // join
}

// ---- static graph of the training step (replaces cuda_graph.h:65-153 of the reference, which captures Trainer::training_step) ----
// One instantiated graph per (batch size, buffer addresses, loss scale, plan version): captured on an internal stream (the caller's
// may be the legacy default stream, which cannot be captured), launched on the caller's.  The only values that change from step
// to step are the four Adam scalars of set_adam_step; they are patched into the graph's optimizer kernel nodes before every launch
// (cudaGraphExecKernelNodeSetParams).  Pointer sets are cached least-recently-used; a caller that passes new buffers every step
// stops being captured after GRAPH_MAX_MISSES misses in a row.
struct StepKey {
	uint32_t B; float loss_scale; uint64_t version;
	const void* ptr[11];
	bool operator==(const StepKey& o) const { return B == o.B && loss_scale == o.loss_scale && version == o.version && !memcmp(ptr, o.ptr, sizeof(ptr)); }
};
struct StepGraph {
	StepKey key{};
	cudaGraph_t graph = nullptr;
	cudaGraphExec_t exec = nullptr;
	uint64_t last_use = 0;
	struct OptimizerNode {
		cudaGraphNode_t node; cudaKernelNodeParams params;
		std::vector<std::vector<unsigned char>> args; std::vector<void*> arg_ptrs; size_t adam_arg;
	};
	std::vector<OptimizerNode> optimizer_nodes;
	~StepGraph() { if (exec) cudaGraphExecDestroy(exec); if (graph) cudaGraphDestroy(graph); }
};
static constexpr size_t GRAPH_CACHE = 8;
static constexpr uint32_t GRAPH_MAX_MISSES = 32;

static bool graphs_enabled(const tcnn_b200_plan* plan) {
	if (plan->graph_broken || plan->profiling) return false;
	if (plan->graph_mode >= 0) return plan->graph_mode != 0;
	static const bool env_on = [] { const char* e = getenv("TCNN_B200_GRAPH"); return !(e && e[0] == '0'); }();
	return env_on;
}

static void drop_step_graphs(tcnn_b200_plan* p) {
	for (StepGraph* g : p->step_graphs) delete g;
	p->step_graphs.clear();
}

// copies the arguments of the optimizer kernels out of a captured graph so that they can be re-issued with new Adam scalars
static void collect_optimizer_nodes(StepGraph& sg) {
	size_t n = 0;
	CUDA_OK(cudaGraphGetNodes(sg.graph, nullptr, &n));
	std::vector<cudaGraphNode_t> nodes(n);
	CUDA_OK(cudaGraphGetNodes(sg.graph, nodes.data(), &n));
	for (cudaGraphNode_t node : nodes) {
		cudaGraphNodeType type;
		CUDA_OK(cudaGraphNodeGetType(node, &type));
		if (type != cudaGraphNodeTypeKernel) continue;
		cudaKernelNodeParams kp{};
		CUDA_OK(cudaGraphKernelNodeGetParams(node, &kp));
		std::vector<size_t> sizes; size_t adam_arg;
		if (kp.func == (void*)adam_kernel) { sizes = {sizeof(AdamArgs)}; adam_arg = 0; }
		else if (kp.func == (void*)mlp_finalize_adam_kernel) { sizes = {4, 4, sizeof(void*), sizeof(void*), sizeof(void*), sizeof(AdamArgs), 4}; adam_arg = 5; }
		else continue;
		CHECK(kp.kernelParams != nullptr, "graph kernel node without a parameter array");
		sg.optimizer_nodes.emplace_back();
		StepGraph::OptimizerNode& on = sg.optimizer_nodes.back();
		on.node = node; on.params = kp; on.adam_arg = adam_arg;
		for (size_t k = 0; k < sizes.size(); ++k) {
			const unsigned char* src = (const unsigned char*)kp.kernelParams[k];
			on.args.emplace_back(src, src + sizes[k]);
		}
		for (auto& a : on.args) on.arg_ptrs.push_back(a.data());
		on.params.kernelParams = on.arg_ptrs.data();
		on.params.extra = nullptr;
	}
}

extern "C" int tcnn_b200_training_step(const tcnn_b200_plan* plan, void* stream_, uint32_t B, const float* x, const float* target,
	float loss_scale, uint32_t adam_step_count, float* params_fp32, void* params_fp16, void* grads_fp16, float* first_moments, float* second_moments,
	uint32_t* param_steps, float* loss_out, void* output_fp16, void* workspace) {
	API_BEGIN
	require_device(plan); check_batch(B);
	CHECK(plan->kind == PLAN_FULL, "training entry points need an encoding and a network");
	CHECK(x && target && params_fp32 && params_fp16 && grads_fp16 && first_moments && second_moments && param_steps && workspace, "null argument");
	ensure_aux(plan);
	cudaStream_t main = (cudaStream_t)stream_;
	tcnn_b200_plan* mp = const_cast<tcnn_b200_plan*>(plan);
	auto direct = [&] {
		training_step_launches(plan, main, B, x, target, loss_scale, adam_step_count, params_fp32, params_fp16, grads_fp16, first_moments, second_moments,
			param_steps, loss_out, output_fp16, workspace);
	};
	cudaStreamCaptureStatus capturing = cudaStreamCaptureStatusNone;
	if (main != nullptr && main != cudaStreamLegacy) CUDA_OK(cudaStreamIsCapturing(main, &capturing));
	if (!graphs_enabled(plan) || capturing != cudaStreamCaptureStatusNone) { direct(); return 0; } // a caller capturing us gets the plain launches

	StepKey key{};
	key.B = B; key.loss_scale = loss_scale; key.version = plan->version;
	const void* ptrs[11] = {x, target, params_fp32, params_fp16, grads_fp16, first_moments, second_moments, param_steps, loss_out, output_fp16, workspace};
	memcpy(key.ptr, ptrs, sizeof(ptrs));
	StepGraph* sg = nullptr;
	for (StepGraph* g : mp->step_graphs) if (g->key == key) { sg = g; break; }
	if (!sg) {
		if (++mp->graph_misses_in_a_row > GRAPH_MAX_MISSES) { // new buffers every step: capturing costs more than it saves
			mp->graph_broken = true; drop_step_graphs(mp);
			direct(); return 0;
		}
		if (mp->step_graphs.size() >= GRAPH_CACHE) {
			size_t oldest = 0;
			for (size_t k = 1; k < mp->step_graphs.size(); ++k) if (mp->step_graphs[k]->last_use < mp->step_graphs[oldest]->last_use) oldest = k;
			delete mp->step_graphs[oldest];
			mp->step_graphs.erase(mp->step_graphs.begin() + oldest);
		}
		if (!mp->capture_stream) CUDA_OK(cudaStreamCreateWithFlags(&mp->capture_stream, cudaStreamNonBlocking));
		std::unique_ptr<StepGraph> fresh(new StepGraph);
		fresh->key = key;
		bool ok = true; std::string why;
		CUDA_OK(cudaStreamBeginCapture(mp->capture_stream, cudaStreamCaptureModeRelaxed));
		try {
			training_step_launches(plan, mp->capture_stream, B, x, target, loss_scale, adam_step_count, params_fp32, params_fp16, grads_fp16, first_moments,
				second_moments, param_steps, loss_out, output_fp16, workspace);
		} catch (const std::exception& e) { ok = false; why = e.what(); }
		const cudaError_t end = cudaStreamEndCapture(mp->capture_stream, &fresh->graph);
		if (ok && end != cudaSuccess) { ok = false; why = std::string("cudaStreamEndCapture: ") + cudaGetErrorString(end); }
		if (ok) {
			const cudaError_t inst = cudaGraphInstantiate(&fresh->exec, fresh->graph, 0);
			if (inst != cudaSuccess) { ok = false; why = std::string("cudaGraphInstantiate: ") + cudaGetErrorString(inst); }
		}
		if (ok) {
			collect_optimizer_nodes(*fresh);
			if (fresh->optimizer_nodes.empty()) { ok = false; why = "no optimizer kernel node found in the captured step"; }
		}
		if (!ok) { // never silently wrong: forget graphs for this plan and say why once
			cudaGetLastError();
			mp->graph_broken = true; drop_step_graphs(mp);
			fprintf(stderr, "tcnn_b200: training-step graph disabled (%s)\n", why.c_str());
			direct(); return 0;
		}
		sg = fresh.release();
		mp->step_graphs.push_back(sg);
		++mp->graph_captures;
	} else {
		mp->graph_misses_in_a_row = 0;
	}
	for (StepGraph::OptimizerNode& on : sg->optimizer_nodes) {
		AdamArgs a;
		memcpy(&a, on.args[on.adam_arg].data(), sizeof(AdamArgs));
		set_adam_step(plan, a, adam_step_count);
		memcpy(on.args[on.adam_arg].data(), &a, sizeof(AdamArgs));
		CUDA_OK(cudaGraphExecKernelNodeSetParams(sg->exec, on.node, &on.params));
	}
	sg->last_use = ++mp->graph_clock;
	CUDA_OK(cudaGraphLaunch(sg->exec, main));
	++mp->graph_replays;
	API_END
}

// mode: 0 = plain launches, 1 = static graphs, -1 = environment TCNN_B200_GRAPH (default: graphs)
extern "C" int tcnn_b200_plan_set_graph_mode(tcnn_b200_plan* plan, int mode) {
	API_BEGIN
	CHECK(plan && mode >= -1 && mode <= 1, "bad argument");
	plan->graph_mode = mode;
	plan->graph_broken = false; plan->graph_misses_in_a_row = 0;
	drop_step_graphs(plan);
	API_END
}
extern "C" int tcnn_b200_plan_graph_stats(const tcnn_b200_plan* plan, uint32_t* captures, uint32_t* replays) {
	API_BEGIN
	CHECK(plan && captures && replays, "null argument");
	*captures = plan->graph_captures; *replays = plan->graph_replays;
	API_END
}


// ---- batches that live in host memory ------------------------------------------------------------------------------------
// The reference's hosts generate or load their batches on the device; a host that holds them in (pinned) host memory needs the
// copy of step i+1 under the compute of step i and the loss back without a device-wide sync.  begin/end wrap ANY of the step
// entry points: begin stages this step's batch (unless an earlier begin already prefetched it), starts the copy of the next one
// into the other slot, makes `stream` wait for the data and hands out device pointers; end releases the slot and sends the
// 4-byte loss to the host on a third stream.  Nothing is allocated: the caller passes `staging` (tcnn_b200_host_staging_bytes).
static size_t host_slot_bytes(const tcnn_b200_plan* plan, uint32_t B) {
	return align256((size_t)B * plan->info.n_input_dims * 4) + align256((size_t)B * plan->info.n_output_dims * 4) + 256;
}
extern "C" size_t tcnn_b200_host_staging_bytes(const tcnn_b200_plan* plan, uint32_t B) { return plan ? 2 * host_slot_bytes(plan, B) : 0; }

static void ensure_pipe(tcnn_b200_plan* p) {
	if (p->pipe.ready) return;
	CUDA_OK(cudaStreamCreateWithFlags(&p->pipe.h2d, cudaStreamNonBlocking));
	CUDA_OK(cudaStreamCreateWithFlags(&p->pipe.d2h, cudaStreamNonBlocking));
	for (int k = 0; k < 2; ++k) {
		CUDA_OK(cudaEventCreateWithFlags(&p->pipe.filled[k], cudaEventDisableTiming));
		CUDA_OK(cudaEventCreateWithFlags(&p->pipe.consumed[k], cudaEventDisableTiming));
		CUDA_OK(cudaEventCreateWithFlags(&p->pipe.loss_written[k], cudaEventDisableTiming));
		CUDA_OK(cudaEventCreateWithFlags(&p->pipe.loss_copied[k], cudaEventDisableTiming));
	}
	p->pipe.ready = true;
}
static void pipe_fill(tcnn_b200_plan* p, int k, uint32_t B, const float* hx, const float* ht) {
	auto& q = p->pipe;
	const size_t xb = (size_t)B * p->info.n_input_dims * 4, tb = (size_t)B * p->info.n_output_dims * 4;
	char* slot = (char*)q.staging[k];
	CUDA_OK(cudaStreamWaitEvent(q.h2d, q.consumed[k], 0));        // the step that last read this slot has finished
	CUDA_OK(cudaMemcpyAsync(slot, hx, xb, cudaMemcpyHostToDevice, q.h2d));
	CUDA_OK(cudaMemcpyAsync(slot + align256(xb), ht, tb, cudaMemcpyHostToDevice, q.h2d));
	CUDA_OK(cudaEventRecord(q.filled[k], q.h2d));
	q.holds_x[k] = hx; q.holds_t[k] = ht; q.holds_B[k] = B;
}

extern "C" int tcnn_b200_host_batch_begin(tcnn_b200_plan* plan, void* stream_, uint32_t B, const float* host_x, const float* host_target,
	const float* next_host_x, const float* next_host_target, void* staging, float** device_x, float** device_target, float** device_loss) {
	API_BEGIN
	CHECK(plan && host_x && host_target && staging && device_x && device_target && device_loss, "null argument");
	require_device(plan); check_batch(B);
	CHECK(!plan->pipe.open, "tcnn_b200_host_batch_begin called twice without tcnn_b200_host_batch_end");
	CHECK((next_host_x == nullptr) == (next_host_target == nullptr), "next batch needs both pointers");
	ensure_pipe(plan);
	auto& q = plan->pipe;
	cudaStream_t main = (cudaStream_t)stream_;
	const size_t slot_bytes = host_slot_bytes(plan, B);
	if (q.staging[0] != staging || q.staging[1] != (char*)staging + slot_bytes) { // new staging buffer (or batch size): nothing is prefetched
		q.staging[0] = staging; q.staging[1] = (char*)staging + slot_bytes;
		q.holds_x[0] = q.holds_x[1] = nullptr;
	}
	// the slot that already holds this batch (prefetched by the previous begin), else the one not used by the previous step
	int k = -1;
	for (int c = 0; c < 2; ++c) if (q.holds_x[c] == host_x && q.holds_t[c] == host_target && q.holds_B[c] == B) k = c;
	if (k < 0) { k = (int)(q.slot ^ 1u); pipe_fill(plan, k, B, host_x, host_target); }
	if (next_host_x) pipe_fill(plan, k ^ 1, B, next_host_x, next_host_target);
	const size_t xb = (size_t)B * plan->info.n_input_dims * 4, tb = (size_t)B * plan->info.n_output_dims * 4;
	char* slot = (char*)q.staging[k];
	CUDA_OK(cudaStreamWaitEvent(main, q.filled[k], 0));
	CUDA_OK(cudaStreamWaitEvent(main, q.loss_copied[k], 0));        // the loss word of this slot (two steps ago) has left the device
	*device_x = (float*)slot; *device_target = (float*)(slot + align256(xb)); *device_loss = (float*)(slot + align256(xb) + align256(tb));
	q.slot = (uint32_t)k; q.open = true;
	q.holds_x[k] = nullptr;                                          // consumed by this step: a later begin with the same host pointer copies again
	API_END
}

extern "C" int tcnn_b200_host_batch_end(tcnn_b200_plan* plan, void* stream_, float* host_loss_out, int join) {
	API_BEGIN
	CHECK(plan && plan->pipe.open, "tcnn_b200_host_batch_end without tcnn_b200_host_batch_begin");
	auto& q = plan->pipe;
	cudaStream_t main = (cudaStream_t)stream_;
	const int k = (int)q.slot;
	q.open = false;
	CUDA_OK(cudaEventRecord(q.consumed[k], main));
	if (host_loss_out) {
		const size_t xb = (size_t)q.holds_B[k] * plan->info.n_input_dims * 4, tb = (size_t)q.holds_B[k] * plan->info.n_output_dims * 4;
		const float* dev_loss = (const float*)((char*)q.staging[k] + align256(xb) + align256(tb));
		CUDA_OK(cudaEventRecord(q.loss_written[k], main));
		CUDA_OK(cudaStreamWaitEvent(q.d2h, q.loss_written[k], 0));
		CUDA_OK(cudaMemcpyAsync(host_loss_out, dev_loss, 4, cudaMemcpyDeviceToHost, q.d2h));
		CUDA_OK(cudaEventRecord(q.loss_copied[k], q.d2h));
		if (join) CUDA_OK(cudaStreamWaitEvent(main, q.loss_copied[k], 0)); // the read-back ends before anything the caller enqueues next
	}
	API_END
}

// ---- data parallel: sharded optimizer over peer memory ---------------------------------------------------------
static void dp_shard(const tcnn_b200_plan* plan, uint32_t rank, uint32_t world, uint64_t* begin, uint64_t* end) {
	const uint64_t n = plan->info.n_params;
	const uint64_t per = ((n + world - 1) / world + 1023) / 1024 * 1024;
	*begin = std::min<uint64_t>((uint64_t)rank * per, n);
	*end = std::min<uint64_t>(*begin + per, n);
}
extern "C" int tcnn_b200_dp_shard(const tcnn_b200_plan* plan, uint32_t rank, uint32_t world_size, uint64_t* param_begin, uint64_t* param_end) {
	API_BEGIN
	CHECK(plan && param_begin && param_end && world_size >= 1 && world_size <= DP_MAX_RANKS && rank < world_size, "bad argument");
	dp_shard(plan, rank, world_size, param_begin, param_end);
	API_END
}
extern "C" uint32_t tcnn_b200_dp_flag_bytes(void) { return DP_FLAG_WORDS * 4; }
extern "C" uint32_t tcnn_b200_dp_flag_error_word(void) { return DP_FLAG_ERROR; }

extern "C" int tcnn_b200_training_step_dp(const tcnn_b200_plan* plan, void* stream_, uint32_t B, const float* x, const float* target,
	float loss_scale, uint32_t adam_step_count, float* params_fp32_shard, float* first_moments_shard, float* second_moments_shard,
	uint32_t* param_steps_shard, float* loss_out, void* output_fp16, void* workspace, const tcnn_b200_dp_comm* comm) {
	API_BEGIN
	require_device(plan); check_batch(B);
	CHECK(plan->kind == PLAN_FULL, "training entry points need an encoding and a network");
	CHECK(x && target && params_fp32_shard && first_moments_shard && second_moments_shard && param_steps_shard && workspace && comm, "null argument");
	CHECK(comm->world_size >= 1 && comm->world_size <= DP_MAX_RANKS && comm->rank < comm->world_size, "bad rank / world size (at most 8 ranks: one NVSwitch domain)");
	CHECK(comm->epoch >= 1, "epoch starts at 1 and grows by one per call");
	const tcnn_b200_info& in = plan->info;
	CHECK(in.n_params % 4 == 0, "parameter count must be a multiple of 4");
	CHECK(plan->enc_kind == ENC_GRID, "the data-parallel step is built for the grid encoding (a parameter-free encoding has ~100 KB of gradients: all-reduce them)");
	DpComm c{};
	c.rank = comm->rank; c.world = comm->world_size; c.epoch = comm->epoch;
	{ // how long a rank waits for its peers inside the kernel before it gives up (sets the error word, does not trap)
		static const double timeout_s = [] { const char* e = getenv("TCNN_B200_DP_TIMEOUT_S"); const double v = e ? atof(e) : 0.0; return v > 0.0 ? v : 600.0; }();
		c.timeout_ns = (uint64_t)(timeout_s * 1e9);
	}
	for (uint32_t r = 0; r < c.world; ++r) {
		CHECK(comm->peer_grads[r] && comm->peer_params[r] && comm->peer_flags[r], "null peer buffer");
		c.peer_grads[r] = (__half*)comm->peer_grads[r]; c.peer_params[r] = (__half*)comm->peer_params[r]; c.peer_flags[r] = (uint32_t*)comm->peer_flags[r];
	}
	c.use_multicast = comm->grads_multicast && comm->params_multicast;
#ifdef DP_STAMPS
	c.use_multicast |= comm->reserved;
#endif
	c.grads_mc = (__half*)comm->grads_multicast; c.params_mc = (__half*)comm->params_multicast;

	ensure_aux(plan);
	cudaStream_t main = (cudaStream_t)stream_, aux = plan->aux;
	const cudaEvent_t* ev = plan->ev;
	const Workspace w = carve(plan, B, workspace);
	const __half* p = c.peer_params[c.rank];
	__half* g = c.peer_grads[c.rank];
	__half* gg = g + in.n_params_mlp;
	const GridParams gp = backward_params(plan, B, true);
	MlpArgs a = base_mlp_args(plan, B, w.encoded, p, &w);
	a.target = target; a.output = (__half*)output_fp16; a.dL_dinput = w.dL_denc;
	a.dW_partials = w.dw_partials; a.loss_partials = w.loss_partials; a.loss_scale = loss_scale;
	const uint32_t mlp_grid = mlp_grid_size(plan, B / MLP_TILE, true);

	// local part: identical to the single-GPU step up to the complete local gradient (grid gradients are zero on entry)
	encode_forward(plan, main, B, x, p + in.n_params_mlp, w.encoded, w.replicas, replica_bytes_of(plan, gp), is_wide(plan)); // also clears the replicas
	if (fused_scatter_available(plan)) {
		a.dL_dinput = nullptr; a.positions = x; a.grid_grad = gg; a.replicas = w.replicas;
		launch_mlp_train_fused(plan, main, a, mlp_grid, gp);
		CUDA_OK(cudaEventRecord(ev[0], main));
		CUDA_OK(cudaStreamWaitEvent(aux, ev[0], 0));
		mlp_finalize(plan, aux, mlp_grid, w, g, loss_out, loss_out != nullptr, B / MLP_TILE); // beside the replica reduction
		CUDA_OK(cudaEventRecord(ev[1], aux));
		backward_finish(plan, main, gp, gg, w.replicas);
		CUDA_OK(cudaStreamWaitEvent(main, ev[1], 0));
	} else {
		launch_mlp<MODE_TRAIN>(plan, main, a, mlp_grid);
		mlp_finalize(plan, main, mlp_grid, w, g, loss_out, loss_out != nullptr, B / MLP_TILE);
		backward_scatter(plan, main, gp, B, x, w.dL_denc, gg, w.replicas);
		backward_finish(plan, main, gp, gg, w.replicas);
	}
	// exchange + optimizer: one kernel (optim_kernels.cuh, dp_adam_kernel)
	uint64_t sb = 0, se = 0;
	dp_shard(plan, c.rank, c.world, &sb, &se);
#ifdef DP_STAMPS
	if (const char* fw = getenv("DP_FAKE_WORLD")) dp_shard(plan, (uint32_t)atoi(getenv("DP_FAKE_RANK")), (uint32_t)atoi(fw), &sb, &se);
#endif
	AdamArgs ad = adam_args(plan, loss_scale * (float)c.world, adam_step_count,
		AdamBuffers{params_fp32_shard - sb, nullptr, nullptr, first_moments_shard - sb, second_moments_shard - sb, param_steps_shard - sb}, 0, (uint32_t)sb, (uint32_t)se);
	{
		ProfileScope prof(plan, main, PK_ADAM);
		const uint32_t n_groups = (uint32_t)((se - sb + 3) / 4);
		const uint32_t maxw = c.world <= 2 ? 2 : c.world <= 4 ? 4 : 8;
		// persistent: as many CTAs as are resident at once walk the shard; an empty shard still takes part in both barriers
		using DpKernel = void (*)(const AdamArgs, const DpComm);
		const DpKernel kernel = (c.use_multicast & 1) ? dp_adam_kernel<true, 8> : maxw == 2 ? dp_adam_kernel<false, 2> : maxw == 4 ? dp_adam_kernel<false, 4> : dp_adam_kernel<false, 8>;
		int per_sm = 0;
		CUDA_OK(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, kernel, 256, 0));
		const uint32_t blocks = std::max(1u, std::min((n_groups + 255) / 256, (uint32_t)plan->n_sms * (uint32_t)std::max(per_sm, 1)));
		kernel<<<blocks, 256, 0, main>>>(ad, c);
		launch_check("dp_adam_kernel");
		// every rank has consumed this rank's gradients once the kernel has retired
		CUDA_OK(cudaMemsetAsync(gg, 0, in.n_params_grid * 2, main));
	}
	API_END
}

// ---- per-kernel timing --------------------------------------------------------------------------------------
extern "C" int tcnn_b200_profile_enable(tcnn_b200_plan* plan, int enable) {
	API_BEGIN
	CHECK(plan, "null plan");
	plan->profiling = enable != 0;
	API_END
}
// Synchronises the recorded events; ms_sum[k] / count[k] for k = grid_fwd, mlp, finalize, grid_bwd, adam; clears the log.
extern "C" int tcnn_b200_profile_read(tcnn_b200_plan* plan, float* ms_sum, uint32_t* count, uint32_t n_kinds) {
	API_BEGIN
	CHECK(plan && ms_sum && count && n_kinds >= (uint32_t)PK_COUNT, "bad argument");
	for (uint32_t k = 0; k < n_kinds; ++k) { ms_sum[k] = 0.f; count[k] = 0; }
	for (auto& sp : plan->spans) {
		CUDA_OK(cudaEventSynchronize(sp.e1));
		float ms = 0.f;
		CUDA_OK(cudaEventElapsedTime(&ms, sp.e0, sp.e1));
		ms_sum[sp.kind] += ms; count[sp.kind] += 1;
		cudaEventDestroy(sp.e0); cudaEventDestroy(sp.e1);
	}
	plan->spans.clear();
	API_END
}
