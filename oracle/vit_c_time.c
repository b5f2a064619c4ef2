/* TEST / BASELINE INFRASTRUCTURE ONLY — never linked into or called from the product path.
 *
 * Timing driver around the reference's C prototype, compiled from the reference source WHERE IT LIES
 * (the Makefile passes -I<reference>/csrc; nothing is copied).  `vit.c` has its own main() (csrc/vit.c:886) and
 * prints seven matrices per head (csrc/vit.c:309-316), so main is renamed on include and stdout is pointed at
 * /dev/null while vit_forward runs; results go to stderr.
 *
 * What it measures: csrc/vit.c:443 vit_forward = patch embedding + CLS + position add + 2 attention blocks,
 * FORWARD ONLY, a different model from the PyTorch path (CLS first, divide by sqrt, no LayerNorm/MLP/head): it is a
 * reported CPU curiosity (north_star: "the reference's CPU path (PyTorch CPU and csrc/vit.c)"), not an oracle for
 * the training step and not comparable in images/s.
 *
 * usage: vit_c_time <image_size> <channels> <patch> <heads> <hidden_dim> <reps>
 *        vit_c_time kat      -> runs the reference's own main() (the 4x4x1 known-answer case of
 *                               tests/multihead-attention-test.ipynb cell 0) with stdout left alone.
 */
#define main vit_c_main
#include "vit.c"
#undef main

#include <unistd.h>
#include <fcntl.h>

static double now_s(void) {
    struct timespec ts;
    clock_gettime(CLOCK_MONOTONIC, &ts);
    return (double)ts.tv_sec + 1e-9 * (double)ts.tv_nsec;
}

int main(int argc, char** argv) {
    if (argc >= 2 && strcmp(argv[1], "kat") == 0) return vit_c_main();
    if (argc < 7) {
        fprintf(stderr, "usage: %s <image_size> <channels> <patch> <heads> <hidden_dim> <reps> | kat\n", argv[0]);
        return 2;
    }
    ViTConfig config;
    memset(&config, 0, sizeof config);
    config.image_size = atoi(argv[1]);
    config.channels = atoi(argv[2]);
    config.patch_size = atoi(argv[3]);
    config.num_layers = 1;
    config.num_heads = atoi(argv[4]);
    config.num_blocks = 2; /* attention_forward is called with 2 blocks, csrc/vit.c:478 */
    config.hidden_dim = atoi(argv[5]);
    config.num_classes = 2;
    int reps = atoi(argv[6]);
    if (config.hidden_dim % config.num_heads != 0 || reps < 1) {
        fprintf(stderr, "bad arguments\n");
        return 2;
    }

    ViT model;
    model.config = config;
    fflush(stdout);
    int saved = dup(1);
    int devnull = open("/dev/null", O_WRONLY);
    dup2(devnull, 1);

    vit_alloc(&model, 0);
    vit_init(&model, 0);
    size_t n_img = (size_t)config.channels * config.image_size * config.image_size;
    float* image = (float*)malloc(n_img * sizeof(float));
    if (!image) return 1;
    initialize_random(image, n_img);

    double best = 1e30, total = 0.0;
    for (int r = 0; r < reps; ++r) {
        double t0 = now_s();
        float* out = vit_forward(&model, image);
        double dt = now_s() - t0;
        free(out);
        total += dt;
        if (dt < best) best = dt;
    }
    fflush(stdout);
    dup2(saved, 1);
    close(devnull);
    close(saved);

    int threads = 1;
#ifdef OMP
    threads = omp_get_max_threads();
#endif
    fprintf(stderr,
            "{\"what\": \"csrc/vit.c vit_forward (patch embed + 2 attention blocks, forward only)\", "
            "\"image_size\": %d, \"channels\": %d, \"patch\": %d, \"heads\": %d, \"hidden_dim\": %d, "
            "\"reps\": %d, \"threads\": %d, \"seconds_per_forward_best\": %.6f, \"seconds_per_forward_mean\": %.6f}\n",
            config.image_size, config.channels, config.patch_size, config.num_heads, config.hidden_dim, reps, threads,
            best, total / reps);
    free(image);
    free(model.params_memory);
    return 0;
}
