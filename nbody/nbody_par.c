/* nbody_par.c -- host driver of the B200 N-body step; builds nbody/nbody-par and nbody/nbody-seq.
 *
 * Written from scratch against include/nbody_gpu.h.  It keeps the observable surface of the
 * reference programs so that bin/nbody-sanity-check and bin/tests.sh see the same thing:
 *   - command line  num_bodies secs_per_update ppm_output_file steps [--run-xps] [--openmp]
 *     [--measure-comm]            (reference nbody/nbody-par.c:816-862, nbody-seq.c:406-435)
 *   - the world size comes from the P6 header of ppm_output_file (nbody-seq.c:239-304)
 *   - body initialisation: srand(27102015), four rand() draws per body (nbody-par.c:208-222)
 *   - stdout: the final state, six "%10.3f" columns per body (nbody-par.c:740-750), or one CSV
 *     line with --run-xps (nbody-par.c:954-957 / nbody-seq.c:488)
 *   - stderr: the banner, "N-body took: %.3f seconds", "Performance N-body: %.2f GFLOPS"
 *     (nbody-par.c:866,943-944); the timer brackets the step loop only (nbody-par.c:895-927)
 * What differs, on purpose: no MAXBODIES cap (bodies are heap arrays), the radius formula's
 * b*b and N*N are formed in double (the reference's int products overflow for N >= 46341),
 * the flop count is 64-bit, and the step loop runs on the GPU(s) behind the C ABI -- there is
 * no CPU step loop in this program.  Extra options, all after the four positionals so the
 * reference's parser would ignore them too: --gpus K (or env NBODY_GPUS), --mode exact|fast
 * (or env NBODY_MODE; default exact), --hex (print the state with %a instead of %10.3f),
 * --display-final (paint the final state into the PPM).  secs_per_update > 0 restores the timed
 * display() of nbody-seq (nbody-seq.c:467-471; dead code in nbody-par), rendered by a CUDA kernel.
 */
#include <dirent.h>
#include <errno.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>
#include <math.h>
#include <sys/time.h>
#include <time.h>

#include "nbody_gpu.h"

#define SEED 27102015

static int skip_space(FILE *f)
{
    int c = fgetc(f);
    while (c == ' ' || c == '\t' || c == '\n' || c == '\r' || c == '#') {
        if (c == '#') {
            while (c != '\n' && c != EOF) c = fgetc(f);
        }
        c = fgetc(f);
    }
    return c;
}

static int read_number(FILE *f, int *out)
{
    int c = skip_space(f);
    if (c < '0' || c > '9') { errno = EPROTO; return -1; }
    int v = 0;
    while (c >= '0' && c <= '9') { v = v * 10 + (c - '0'); c = fgetc(f); }
    if (c != EOF) ungetc(c, f);
    *out = v;
    return 0;
}

/* Width and height from a binary PPM header: "P6" <w> <h> <maxval<=255> <one whitespace>. */
static int read_p6_dims(const char *path, int *xdim, int *ydim, long *data_offset)
{
    FILE *f = fopen(path, "rb");
    int maxval = 0;
    if (!f) return -1;
    if (fgetc(f) != 'P' || fgetc(f) != '6') { errno = EPROTO; fclose(f); return -1; }
    if (read_number(f, xdim) || read_number(f, ydim) || read_number(f, &maxval) || maxval > 255) {
        if (!errno) errno = EPROTO;
        fclose(f);
        return -1;
    }
    {
        int c = fgetc(f);   /* exactly one whitespace character, then the 24-bit data (nbody-seq.c:287-296) */
        if (c != ' ' && c != '\t' && c != '\n' && c != '\r') { errno = EPROTO; fclose(f); return -1; }
    }
    *data_offset = ftell(f);
    fclose(f);
    return 0;
}

/* display() + msync of the reference (nbody-seq.c:467-471): paint the current state into the P6 file. */
static void write_image(nbody_gpu *h, const char *path, long data_offset, int xdim, int ydim)
{
    size_t bytes = (size_t)xdim * ydim * 3;
    unsigned char *img = malloc(bytes);
    FILE *f;
    if (!img) {
        fprintf(stderr, "Cannot malloc(%zu) for the image of %s\n", bytes, path);
        exit(1);
    }
    if (nbody_gpu_render(h, img) != 0) {
        fprintf(stderr, "nbody_gpu_render: %s\n", nbody_gpu_last_error());
        exit(1);
    }
    f = fopen(path, "r+b");
    if (f) {
        if (fseek(f, data_offset, SEEK_SET) == 0) fwrite(img, 1, bytes, f);
        fclose(f);
    } else {
        fprintf(stderr, "Cannot write %s: %s\n", path, strerror(errno));
    }
    free(img);
}

#ifndef NBODY_SEQ
/* --run-xps: the reference reads NODES and CPUS_PER_NODE back from the <job>.PRUN_ENVIRONMENT file the cluster
 * launcher leaves in the working directory (nbody-par.c:441-517: PRUN_CPUS / PRUN_CPUS_PER_NODE and
 * PRUN_CPUS_PER_NODE).  Same here (bin/prun writes that file); without one: 1 node, `fallback_cpus` per node. */
static void prun_compute_args(int *nodes, int *cpus_per_node, int fallback_cpus)
{
    DIR *dr = opendir(".");
    struct dirent *en;
    char name[512] = "";
    *nodes = 1;
    *cpus_per_node = fallback_cpus;
    if (!dr) return;
    while ((en = readdir(dr)) != NULL) {
        if (strstr(en->d_name, ".PRUN_ENVIRONMENT") != NULL) {
            snprintf(name, sizeof name, "%s", en->d_name);
            break;
        }
    }
    closedir(dr);
    if (!name[0]) return;
    FILE *f = fopen(name, "r");
    if (!f) return;
    char line[1000];
    int cpus = 0, per_node = 0;
    while (fgets(line, sizeof line, f) != NULL) {
        if (strncmp(line, "PRUN_CPUS=", 10) == 0) cpus = atoi(line + 10);
        else if (strncmp(line, "PRUN_CPUS_PER_NODE=", 19) == 0) per_node = atoi(line + 19);
    }
    fclose(f);
    if (cpus > 0 && per_node > 0) {
        *nodes = cpus / per_node;   /* prun multiplies -np by -<cpus>: dividing gives back the -np argument */
        *cpus_per_node = per_node;
    }
}
#endif

static long long nr_flops(long long n, long long steps)
{
    /* the reference's model, nbody-par.c:752-765 */
    return steps * (20 * (n * (n - 1) / 2) + 18 * n + 4 * n);
}

static double now_seconds(void)
{
    struct timeval tv;
    gettimeofday(&tv, 0);
    return tv.tv_sec + tv.tv_usec / 1000000.0;
}

static void die_gpu(const char *what)
{
    fprintf(stderr, "%s: %s\n", what, nbody_gpu_last_error());
    exit(1);
}

int main(int argc, char **argv)
{
    int run_xps = 0, measure_comm = 0, hex = 0, display_final = 0;
    long data_offset = 0;
#ifdef NBODY_SEQ
    int ngpus = 1;
    const char *usage_flags = "[--run-xps]";
#else
    int ngpus = 0;
    const char *usage_flags = "[--run-xps, --openmp, --measure-comm]";
#endif
    int mode = NBODY_GPU_MODE_EXACT;
    const char *env;

    if (argc < 5) {
        fprintf(stderr, "Usage: %s num_bodies secs_per_update ppm_output_file steps %s\n", argv[0], usage_flags);
        exit(1);
    }
    if ((env = getenv("NBODY_GPUS")) != NULL && ngpus == 0) ngpus = atoi(env);
    if ((env = getenv("NBODY_MODE")) != NULL && strcmp(env, "fast") == 0) mode = NBODY_GPU_MODE_FAST;
    for (int i = 5; i < argc; i++) {
        if (strcmp(argv[i], "--run-xps") == 0) run_xps = 1;
        else if (strcmp(argv[i], "--measure-comm") == 0) measure_comm = 1;
        else if (strcmp(argv[i], "--openmp") == 0) { /* accepted: the GPU threads are the worksharing */ }
        else if (strcmp(argv[i], "--hex") == 0) hex = 1;
        else if (strcmp(argv[i], "--display-final") == 0) display_final = 1;
        else if (strcmp(argv[i], "--gpus") == 0 && i + 1 < argc) ngpus = atoi(argv[++i]);
        else if (strcmp(argv[i], "--mode") == 0 && i + 1 < argc)
            mode = strcmp(argv[++i], "fast") == 0 ? NBODY_GPU_MODE_FAST : NBODY_GPU_MODE_EXACT;
        /* anything else is ignored, like the reference does */
    }

    long nb = atol(argv[1]);
    if (nb < 2) {
        fprintf(stderr, "Using two bodies...\n");
        nb = 2;
    }
    if (nb > 0x7fffffffL / 2) nb = 0x7fffffffL / 2;
    const int n = (int)nb;
    const int secsup = atoi(argv[2]);   /* secs_per_update > 0: paint the world into the PPM every secsup seconds */
    int xdim = 0, ydim = 0;
    if (read_p6_dims(argv[3], &xdim, &ydim, &data_offset) == -1) {
        fprintf(stderr, "Cannot read %s: %s\n", argv[3], strerror(errno));
        exit(1);
    }
    const int steps = atoi(argv[4]);

    fprintf(stderr, "Running N-body with %i bodies and %i steps\n", n, steps);

    double *buf = calloc((size_t)n * 8, sizeof(double));
    if (buf == NULL) {
        fprintf(stderr, "Cannot calloc(world)\n");
        exit(1);
    }
    double *x = buf, *y = buf + n, *xf = buf + 2 * (size_t)n, *yf = buf + 3 * (size_t)n;
    double *xv = buf + 4 * (size_t)n, *yv = buf + 5 * (size_t)n, *mass = buf + 6 * (size_t)n, *radius = buf + 7 * (size_t)n;

    srand(SEED);
    const double diag = sqrt(1.0 * ((xdim * xdim) + (ydim * ydim)));
    for (int b = 0; b < n; ++b) {
        x[b] = (rand() % xdim);
        y[b] = (rand() % ydim);
        radius[b] = 1 + (((double)b * b + 1.0) * diag) / (25.0 * ((double)n * n + 1.0));
        mass[b] = radius[b] * radius[b] * radius[b];
        xv[b] = ((rand() % 20000) - 10000) / 2000.0;
        yv[b] = ((rand() % 20000) - 10000) / 2000.0;
    }

    const int visible = nbody_gpu_device_count();
    if (visible < 1) {
        fprintf(stderr, "No CUDA device: this build has no CPU step loop\n");
        exit(1);
    }
    if (ngpus <= 0) ngpus = 1;
    if (ngpus > visible) ngpus = visible;   /* -np larger than the box: same results, fewer shards */
    if (ngpus > n) ngpus = n;

    nbody_gpu_cfg cfg;
    memset(&cfg, 0, sizeof cfg);
    cfg.n = n; cfg.xdim = xdim; cfg.ydim = ydim; cfg.mode = mode;
    cfg.ngpus = ngpus; cfg.device = ngpus > 1 ? -1 : 0;
    nbody_gpu *h = NULL;
    /* the world is allocated once and lives for the whole run (nbody-par.c:810): page-lock it so that the upload
     * and the final fetch run at PCIe speed; the run works without (pageable copies are staged by the driver) */
    const int pinned = nbody_gpu_pin_host(buf, (size_t)n * 8 * sizeof(double)) == 0;
    if (nbody_gpu_init(&h, &cfg, x, y, xv, yv, mass, radius) != 0) die_gpu("nbody_gpu_init");

    double comm_overhead = 0;
    (void)comm_overhead;
    const double start = now_seconds();
    if (measure_comm && run_xps) {
        /* nbody-par.c:912,918: time spent in the exchange, here from device timers between the phases of every
         * step; that needs plain launches (no CUDA graph), so RTIME of such a run includes the launch gaps */
        float ms[4];
        if (nbody_gpu_step_profiled(h, steps, 0, ms) != 0) die_gpu("nbody_gpu_step_profiled");
        comm_overhead = ms[3] / 1000.0;
    } else if (secsup > 0) {
        /* nbody-seq.c:457-471: the step loop with the timed display; the clock is looked at every 16 steps */
        long lastup = 0;
        for (int done = 0; done < steps;) {
            int chunk = steps - done < 16 ? steps - done : 16;
            if (nbody_gpu_step(h, chunk) != 0) die_gpu("nbody_gpu_step");
            done += chunk;
            if ((long)time(0) - lastup > secsup) {
                write_image(h, argv[3], data_offset, xdim, ydim);
                lastup = (long)time(0);
            }
        }
    } else {
        if (nbody_gpu_step(h, steps) != 0) die_gpu("nbody_gpu_step");
    }
    if (nbody_gpu_sync(h) != 0) die_gpu("nbody_gpu_sync");
    const double rtime = now_seconds() - start;

    if (display_final) write_image(h, argv[3], data_offset, xdim, ydim);
    if (nbody_gpu_fetch(h, x, y, xf, yf, xv, yv) != 0) die_gpu("nbody_gpu_fetch");
    const double gflops = nr_flops(n, steps) / 1e9 / rtime;

    if (!run_xps) {
        const char *fmt = hex ? "%a %a %a %a %a %a\n" : "%10.3f %10.3f %10.3f %10.3f %10.3f %10.3f\n";
        for (int b = 0; b < n; ++b) printf(fmt, x[b], y[b], xf[b], yf[b], xv[b], yv[b]);
        fflush(stdout);
    } else {
#ifdef NBODY_SEQ
        printf("%d,%.3f, %.2f\n", n, rtime, gflops);
#else
        /* SIZE,NODES,CPUS_PER_NODE,NBODIES,RTIME,COMMTIME,COMMTIME:RTIME,GFLOPS (nbody-par.c:954-957): SIZE = the
         * GPUs that shared the rows (the reference: MPI ranks), NODES/CPUS_PER_NODE as the launcher recorded them */
        int nodes = 1, cpus_per_node = ngpus;
        prun_compute_args(&nodes, &cpus_per_node, ngpus);
        printf("%d,%d,%d,%d,%.3f,%.3f,%.3f,%.2f\n", ngpus, nodes, cpus_per_node, n, rtime, comm_overhead,
               comm_overhead / rtime, gflops);
#endif
    }
    fprintf(stderr, "\nN-body took: %.3f seconds\n", rtime);
    fprintf(stderr, "Performance N-body: %.2f GFLOPS\n", gflops);
    fprintf(stderr, "GPU: %d x %s mode, %.4g pair-interactions/s, %.4g steps/s\n", ngpus,
            mode == NBODY_GPU_MODE_FAST ? "fast" : "exact",
            rtime > 0 ? (double)steps * n * (n - 1.0) / rtime : 0.0, rtime > 0 ? steps / rtime : 0.0);

    nbody_gpu_destroy(h);
    if (pinned) nbody_gpu_unpin_host(buf);
    free(buf);
    return 0;
}
