// API conformance host: the training loop of the reference's own sample (samples/mlp_learning_an_image.cu:185-315) written against
// include/tiny-cuda-nn/ of THIS repository.  Everything between the two "reference loop" markers uses the reference's public C++ API
// exactly as the sample does - json config, create_loss / create_optimizer, NetworkWithInputEncoding<precision_t>,
// Trainer<float, precision_t, precision_t>, default_rng_t + generate_random_uniform, linear_kernel, GPUMatrix, training_step / loss /
// inference, free_all_gpu_memory_arenas.  Only the image is replaced: there is no image decoder on the box, so the texture
// lookup kernel of the sample (eval_image) evaluates an analytic RGB function of the coordinate instead of a texture.
//
//   nvcc -gencode arch=compute_100a,code=sm_100a -std=c++17 -I include tests/cpp/sample_training_loop.cu -L tiny-cuda-nn_b200/lib -ltcnn_b200
//
// Exit code 0: the loss fell by 20x, inference of the trained network reproduces the target, a snapshot survives
// json::to_msgpack -> from_msgpack -> Trainer::deserialize on a fresh trainer, and the unsupported arguments throw.
#include <tiny-cuda-nn/common_device.h>

#include <tiny-cuda-nn/config.h>

#include <chrono>
#include <cmath>
#include <cstdlib>
#include <fstream>
#include <iomanip>
#include <iterator>
#include <iostream>
#include <random>
#include <stdexcept>
#include <string>
#include <thread>
#include <vector>

using namespace tcnn;
using precision_t = network_precision_t;

// stands in for the sample's eval_image<stride>(n, texture, xs_and_ys, result): same signature shape, analytic "image"
template <uint32_t stride>
__global__ void eval_image(uint32_t n_elements, const float* __restrict__ xs_and_ys, float* __restrict__ result) {
	uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
	if (i >= n_elements) return;
	uint32_t output_idx = i * stride;
	uint32_t input_idx = i * 2;
	const float s = xs_and_ys[input_idx] + 1.5f * xs_and_ys[input_idx + 1];
	for (uint32_t c = 0; c < 3; ++c) result[output_idx + c] = 0.5f + 0.5f * sinf(6.2831853f * (float)(c + 1) * s);
	for (uint32_t c = 3; c < stride; ++c) result[output_idx + c] = 1;
}

// --resume <cfg.json> <snapshot.msgpack> <B> <seed> <n_steps>: Trainer::deserialize of a snapshot the REFERENCE wrote after n_steps of
// `ref_driver serialize` (oracle/ref_driver.cu), then the two steps that follow on the same batch stream; prints their losses.
static int resume(char* argv[]) {
	std::ifstream cf{argv[2]};
	json config = json::parse(cf, nullptr, true, true);
	std::ifstream sf(argv[3], std::ios::binary);
	const std::vector<uint8_t> raw((std::istreambuf_iterator<char>(sf)), std::istreambuf_iterator<char>());
	const uint32_t B = atoi(argv[4]), n_steps = atoi(argv[6]);
	const uint32_t n_in = 2, n_out = 3;
	cudaStream_t stream;
	CUDA_CHECK_THROW(cudaStreamCreate(&stream));
	TrainableModel model = create_from_config(n_in, n_out, config);
	model.trainer->deserialize(json::from_msgpack(raw));
	default_rng_t rng{(uint64_t)atoll(argv[5])};
	GPUMatrix<float> input(n_in, B), target(n_out, B);
	for (uint32_t i = 0; i < n_steps; ++i) generate_random_uniform<float>(stream, rng, B * n_in, input.data());
	std::cout << "resume losses";
	for (int i = 0; i < 2; ++i) {
		generate_random_uniform<float>(stream, rng, B * n_in, input.data());
		linear_kernel(eval_image<n_out>, 0, stream, B, input.data(), target.data());
		auto ctx = model.trainer->training_step(stream, input, target);
		std::cout << " " << std::setprecision(9) << model.trainer->loss(stream, *ctx);
	}
	std::cout << " step " << model.optimizer->step() << std::endl;
	return 0;
}

int main(int argc, char* argv[]) {
	try {
		if (argc >= 7 && std::string(argv[1]) == "--resume") return resume(argv);

		uint32_t compute_capability = cuda_compute_capability();
		if (compute_capability < MIN_GPU_ARCH) {
			std::cerr
				<< "Warning: Insufficient compute capability " << compute_capability << " detected. "
				<< "This program was compiled for >=" << MIN_GPU_ARCH << " and may thus behave unexpectedly." << std::endl;
		}

		json config = {
			{"loss", {
				{"otype", "RelativeL2"}
			}},
			{"optimizer", {
				{"otype", "Adam"},
				// {"otype", "Shampoo"},
				{"learning_rate", 1e-2},
				{"beta1", 0.9f},
				{"beta2", 0.99f},
				{"l2_reg", 0.0f},
				// The following parameters are only used when the optimizer is "Shampoo".
				{"beta3", 0.9f},
				{"beta_shampoo", 0.0f},
				{"identity", 0.0001f},
				{"cg_on_momentum", false},
				{"frobenius_normalization", true},
			}},
			// the sample's inline default is OneBlob + 64x4; this library implements the HashGrid path, so the default here is the
			// reference's data/config_hash.json (the configuration BASELINE.json quotes)
			{"encoding", {
				{"otype", "HashGrid"},
				{"n_levels", 16},
				{"n_features_per_level", 2},
				{"log2_hashmap_size", 15},
				{"base_resolution", 16},
				{"per_level_scale", 1.5},
			}},
			{"network", {
				{"otype", "FullyFusedMLP"},
				// {"otype", "CutlassMLP"},
				{"n_neurons", 64},
				{"n_hidden_layers", 2},
				{"activation", "ReLU"},
				{"output_activation", "None"},
			}},
		};

		if (argc >= 2) {
			std::cout << "Loading custom json config '" << argv[1] << "'." << std::endl;
			std::ifstream f{argv[1]};
			config = json::parse(f, nullptr, true, /*skip_comments=*/true);
		}

		int sampling_width = 256;
		int sampling_height = 256;

		uint32_t n_coords = sampling_width * sampling_height;
		uint32_t n_coords_padded = next_multiple(n_coords, BATCH_SIZE_GRANULARITY);

		GPUMemory<float> sampled_image(n_coords * 3);
		GPUMemory<float> xs_and_ys(n_coords_padded * 2);

		std::vector<float> host_xs_and_ys(n_coords * 2);
		for (int y = 0; y < sampling_height; ++y) {
			for (int x = 0; x < sampling_width; ++x) {
				int idx = (y * sampling_width + x) * 2;
				host_xs_and_ys[idx+0] = (float)(x + 0.5) / (float)sampling_width;
				host_xs_and_ys[idx+1] = (float)(y + 0.5) / (float)sampling_height;
			}
		}

		xs_and_ys.copy_from_host(host_xs_and_ys.data());

		linear_kernel(eval_image<3>, 0, nullptr, n_coords, xs_and_ys.data(), sampled_image.data());

		// ---- reference loop: samples/mlp_learning_an_image.cu:185-296 (texture argument of eval_image dropped, image writes dropped) ----

		// Fourth step: train the model by sampling the above image and optimizing an error metric

		// Various constants for the network and optimization
		const uint32_t batch_size = 1 << 16;
		const uint32_t n_training_steps = argc >= 3 ? atoi(argv[2]) : 300;
		const uint32_t n_input_dims = 2; // 2-D image coordinate
		const uint32_t n_output_dims = 3; // RGB color

		cudaStream_t inference_stream;
		CUDA_CHECK_THROW(cudaStreamCreate(&inference_stream));
		cudaStream_t training_stream = inference_stream;

		default_rng_t rng{1337};

		// Auxiliary matrices for training
		GPUMatrix<float> training_target(n_output_dims, batch_size);
		GPUMatrix<float> training_batch(n_input_dims, batch_size);

		// Auxiliary matrices for evaluation
		GPUMatrix<float> prediction(n_output_dims, n_coords_padded);
		GPUMatrix<float> inference_batch(xs_and_ys.data(), n_input_dims, n_coords_padded);

		json encoding_opts = config.value("encoding", json::object());
		json loss_opts = config.value("loss", json::object());
		json optimizer_opts = config.value("optimizer", json::object());
		json network_opts = config.value("network", json::object());

		std::shared_ptr<Loss<precision_t>> loss{create_loss<precision_t>(loss_opts)};
		std::shared_ptr<Optimizer<precision_t>> optimizer{create_optimizer<precision_t>(optimizer_opts)};
		std::shared_ptr<NetworkWithInputEncoding<precision_t>> network = std::make_shared<NetworkWithInputEncoding<precision_t>>(n_input_dims, n_output_dims, encoding_opts, network_opts);

		auto trainer = std::make_shared<Trainer<float, precision_t, precision_t>>(network, optimizer, loss);

		std::chrono::steady_clock::time_point begin = std::chrono::steady_clock::now();

		float tmp_loss = 0;
		uint32_t tmp_loss_counter = 0;

		std::cout << "Beginning optimization with " << n_training_steps << " training steps." << std::endl;

		uint32_t interval = 10;
		float first_loss = -1, last_loss = -1;

		for (uint32_t i = 0; i < n_training_steps; ++i) {
			bool print_loss = i % interval == 0;
			bool visualize_learned_func = argc < 5 && i % interval == 0;

			// Compute reference values at random coordinates
			{
				generate_random_uniform<float>(training_stream, rng, batch_size * n_input_dims, training_batch.data());
				linear_kernel(eval_image<n_output_dims>, 0, training_stream, batch_size, training_batch.data(), training_target.data());
			}

			// Training step
			{
				auto ctx = trainer->training_step(training_stream, training_batch, training_target);

				if (i % std::min(interval, (uint32_t)100) == 0) {
					tmp_loss += trainer->loss(training_stream, *ctx);
					++tmp_loss_counter;
				}
			}

			// Debug outputs
			{
				if (print_loss) {
					std::chrono::steady_clock::time_point end = std::chrono::steady_clock::now();
					std::cout << "Step#" << i << ": " << "loss=" << tmp_loss/(float)tmp_loss_counter << " time=" << std::chrono::duration_cast<std::chrono::microseconds>(end - begin).count() << "[µs]" << std::endl;
					if (first_loss < 0) first_loss = tmp_loss/(float)tmp_loss_counter;
					last_loss = tmp_loss/(float)tmp_loss_counter;

					tmp_loss = 0;
					tmp_loss_counter = 0;
				}

				if (visualize_learned_func) {
					network->inference(inference_stream, inference_batch, prediction);
				}

				// Don't count visualizing as part of timing
				// (assumes visualize_learned_pdf is only true when print_loss is true)
				if (print_loss) {
					begin = std::chrono::steady_clock::now();
				}
			}

			if (print_loss && i > 0 && interval < 1000) {
				interval *= 10;
			}
		}

		network->inference(inference_stream, inference_batch, prediction);

		// ---- end of the reference loop ----

		CUDA_CHECK_THROW(cudaStreamSynchronize(inference_stream));
		std::vector<float> got(n_coords * 3), want(n_coords * 3);
		CUDA_CHECK_THROW(cudaMemcpy(got.data(), prediction.data(), got.size() * sizeof(float), cudaMemcpyDeviceToHost));
		sampled_image.copy_to_host(want.data());
		double mse = 0;
		for (size_t k = 0; k < got.size(); ++k) mse += (got[k] - want[k]) * (got[k] - want[k]);
		mse /= got.size();
		std::cout << "loss first=" << first_loss << " last=" << last_loss << " inference mse=" << mse << std::endl;

		// Trainer::serialize -> msgpack -> Trainer::deserialize on a fresh model (trainer.h:275-315): same parameters, and the next step
		// of both trainers gives the same loss
		const std::vector<uint8_t> blob = json::to_msgpack(trainer->serialize(true));
		TrainableModel other = create_from_config(n_input_dims, n_output_dims, config);
		other.trainer->deserialize(json::from_msgpack(blob));
		generate_random_uniform<float>(training_stream, rng, batch_size * n_input_dims, training_batch.data());
		linear_kernel(eval_image<n_output_dims>, 0, training_stream, batch_size, training_batch.data(), training_target.data());
		auto ctx_a = trainer->training_step(training_stream, training_batch, training_target);
		const float la = trainer->loss(training_stream, *ctx_a);
		auto ctx_b = other.trainer->training_step(training_stream, training_batch, training_target);
		const float lb = other.trainer->loss(training_stream, *ctx_b);
		std::cout << "resumed: loss " << la << " vs " << lb << " (snapshot " << blob.size() << " bytes, optimizer step " << other.optimizer->step() << ")" << std::endl;

		// hyper-parameter update (trainer.h:213-216) and the arguments the fused step does not implement
		trainer->update_hyperparams({{"optimizer", {{"learning_rate", 1e-3}}}});
		bool threw = false;
		try { GPUMatrix<float> pdf(n_output_dims, batch_size); trainer->training_step(training_stream, training_batch, training_target, &pdf); } catch (const std::runtime_error&) { threw = true; }

		free_all_gpu_memory_arenas();

		const bool ok = last_loss < 0.05f * first_loss && mse < 5e-3 && std::fabs(la - lb) <= 0.05f * std::fabs(la) && threw && std::fabs(trainer->optimizer()->learning_rate() - 1e-3f) < 1e-9f;
		return ok ? EXIT_SUCCESS : EXIT_FAILURE;
	} catch (const std::exception& e) {
		std::cout << "Uncaught exception: " << e.what() << std::endl;
		return 2;
	}
}
