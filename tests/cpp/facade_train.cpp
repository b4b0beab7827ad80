// Host written against the reference's C++ API names (create_from_config, Trainer::training_step, Trainer::loss,
// Network::inference, GPUMatrix) using the facade headers in include/tiny-cuda-nn.  Trains the README config on an
// analytic target and checks that the loss falls and that inference reproduces the trained function.
//   build: g++ -std=c++17 -I include -I /usr/local/cuda/include tests/cpp/facade_train.cpp -L tiny-cuda-nn_b200/lib -ltcnn_b200 -L /usr/local/cuda/lib64 -lcudart
#include <tiny-cuda-nn/config.h>
#include <tiny-cuda-nn/cpp_api.h>

#include <cmath>
#include <cstdio>
#include <fstream>
#include <iterator>
#include <random>
#include <vector>

using namespace tcnn;

static const char* kConfig = R"({
	"loss": {"otype": "L2"},
	"optimizer": {"otype": "Adam", "learning_rate": 1e-3},
	"encoding": {"otype": "HashGrid", "n_levels": 16, "n_features_per_level": 2, "log2_hashmap_size": 19, "base_resolution": 16, "per_level_scale": 2.0},
	"network": {"otype": "FullyFusedMLP", "activation": "ReLU", "output_activation": "None", "n_neurons": 64, "n_hidden_layers": 2}
})";

int main(int argc, char** argv) {
	const bool compile_only = argc > 1 && std::string(argv[1]) == "--no-gpu";
	try {
		if (argc > 2 && std::string(argv[1]) == "--msgpack") { // no device: a snapshot written by the reference goes through json.h unchanged
			std::ifstream f(argv[2], std::ios::binary);
			const std::vector<uint8_t> raw((std::istreambuf_iterator<char>(f)), std::istreambuf_iterator<char>());
			const json snap = json::from_msgpack(raw);
			const bool same = json::to_msgpack(snap) == raw;
			std::printf("n_params=%zu params_type=%s params_bytes=%zu current_step=%u moments_bytes=%zu rewritten_identically=%d\n", (size_t)snap["n_params"],
				snap["params_type"].get<std::string>().c_str(), snap["params_binary"].get_binary().size(), (uint32_t)snap["optimizer"]["current_step"],
				snap["optimizer"]["first_moments_binary"].get_binary().size(), (int)same);
			return same ? 0 : 1;
		}
		const uint32_t n_in = 2, n_out = 3, B = 1 << 14, steps = 300;
		if (compile_only) { // no device here: the plan (JSON, level table, sizes) can still be built; compute calls must throw
			Network net(kConfig, n_in, n_out);
			std::printf("n_params=%zu padded_output_width=%u\n", net.n_params(), net.padded_output_width());
			return net.n_params() == 11191808 && net.padded_output_width() == 16 ? 0 : 1;
		}
		cudaStream_t stream; CUDA_CHECK_THROW(cudaStreamCreate(&stream));
		TrainableModel model = create_from_config(n_in, n_out, kConfig);
		GPUMatrix<float> input(n_in, B), target(n_out, B), prediction(n_out, B);
		std::mt19937 rng(7); std::uniform_real_distribution<float> U(0.f, 1.f);
		std::vector<float> x(B * n_in), t(B * n_out);
		float first = 0.f, last = 0.f;
		for (uint32_t i = 0; i < steps; ++i) {
			for (uint32_t s = 0; s < B; ++s) {
				x[s * 2] = U(rng); x[s * 2 + 1] = U(rng);
				const float ph = x[s * 2] + 1.5f * x[s * 2 + 1];
				for (uint32_t c = 0; c < n_out; ++c) t[s * n_out + c] = 0.5f + 0.5f * std::sin(6.2831853f * (c + 1) * ph);
			}
			input.copy_from_host(x.data()); target.copy_from_host(t.data());
			auto ctx = model.trainer->training_step(stream, input, target);
			const float l = model.trainer->loss(stream, *ctx);
			if (i == 0) first = l;
			last = l;
		}
		model.network->inference(stream, input, prediction);
		CUDA_CHECK_THROW(cudaStreamSynchronize(stream));
		const std::vector<float> p = prediction.to_cpu_vector();
		double mse = 0; for (size_t k = 0; k < p.size(); ++k) mse += (p[k] - t[k]) * (p[k] - t[k]); mse /= p.size();
		std::printf("loss first=%g last=%g inference mse=%g\n", first, last, mse);
		bool threw = false;
		try { GPUMatrix<float> bad(n_in, 300); model.network->inference(stream, bad, prediction); } catch (const std::runtime_error&) { threw = true; }
		return (last < 0.05f * first && mse < 2.0 * last + 1e-4 && threw) ? 0 : 1;
	} catch (const std::exception& e) {
		std::fprintf(stderr, "error: %s\n", e.what());
		return 2;
	}
}
