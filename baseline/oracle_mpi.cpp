// Forest emulation: the reference's kdtree_mpi.cpp translation
// unit (slice generator, build_tree, nearest_neighbor) with `main` renamed and
// mpi.h supplied by baseline/mpi_shim.  The dispatcher below repeats only the
// driver's slice arithmetic so that problem sizes other than the hard-coded
// 500000 x 128 can be run (the reference MPI file does not compile with DEBUG 1).
//
//   oracle_mpi MODE SEED DIM N NQ W [RANK]
//     MODE = lines   per-query min over the W local trees, printed as the reference prints them
//            queryx  "<q> <rank> <nn id> <dist bits hex>" for every rank (or only RANK)
//            time    JSON: per-rank generate/build/query seconds, run in W concurrent processes
#define main kd_reference_mpi_main
#include "_ref/src/kdtree_mpi.cpp"
#undef main

#include <cstdio>
#include <cstring>
#include <string>
#include <vector>

struct RankResult { double gen_s, build_s, query_s; std::vector<float> dist; std::vector<int> ids; };

static RankResult run_rank(int seed, int dim, int n, int nq, int W, int r) {
    RankResult rr;
    int local = n / W, rem = n % W, first = local * r;
    if (r == W - 1) local += rem;
    auto t0 = std::chrono::high_resolution_clock::now();
    float* x = generate_problem(seed, dim, local, first, n, nq);
    Point** pts = (Point**)calloc(local, sizeof(Point*));
    for (int i = 0; i < local; ++i) pts[i] = new Point(dim, i + first + 1, x + (size_t)(i + nq) * dim);
    auto t1 = std::chrono::high_resolution_clock::now();
    Node* tree = build_tree(pts, local);
    auto t2 = std::chrono::high_resolution_clock::now();
    for (int q = 0; q < nq; ++q) {
        Point query(dim, n + q, x + (size_t)q * dim);
        Node* res = nearest_neighbor(tree, &query);
        rr.dist.push_back(query.distance(*res->point));
        rr.ids.push_back(res->point->ID);
    }
    auto t3 = std::chrono::high_resolution_clock::now();
    auto s = [](auto a, auto b) { return std::chrono::duration<double>(b - a).count(); };
    rr.gen_s = s(t0, t1); rr.build_s = s(t1, t2); rr.query_s = s(t2, t3);
    return rr;
}

int main(int argc, char** argv) {
    if (argc < 7) { fprintf(stderr, "usage: %s MODE SEED DIM N NQ W [RANK]\n", argv[0]); return 2; }
    std::string mode = argv[1];
    int seed = atoi(argv[2]), dim = atoi(argv[3]), n = atoi(argv[4]), nq = atoi(argv[5]), W = atoi(argv[6]);
    int only = argc > 7 ? atoi(argv[7]) : -1;
    if (W < 1 || n / W < 1) { fprintf(stderr, "need 1 <= W <= N\n"); return 2; }
    if (mode == "time") {
        // W concurrent processes, like `mpirun -np W`; results come back through pipes
        std::vector<int> fds(W);
        for (int r = 0; r < W; ++r) {
            int p[2]; if (pipe(p)) return 2;
            pid_t c = fork();
            if (c == 0) {
                close(p[0]);
                RankResult rr = run_rank(seed, dim, n, nq, W, r);
                double v[3] = {rr.gen_s, rr.build_s, rr.query_s};
                if (write(p[1], v, sizeof v) != (ssize_t)sizeof v) _exit(3);
                _exit(0);
            }
            close(p[1]); fds[r] = p[0];
        }
        double mx[3] = {0, 0, 0};
        for (int r = 0; r < W; ++r) {
            double v[3] = {0, 0, 0};
            if (read(fds[r], v, sizeof v) != (ssize_t)sizeof v) return 3;
            for (int k = 0; k < 3; ++k) if (v[k] > mx[k]) mx[k] = v[k];
            int st; wait(&st);
        }
        printf("{\"n\": %d, \"dim\": %d, \"nq\": %d, \"world\": %d, \"generate_s\": %.6f, \"build_s\": %.6f, "
               "\"query_s\": %.6f}\n", n, dim, nq, W, mx[0], mx[1], mx[2]);
        return 0;
    }
    std::vector<float> best(nq, 0.f);
    for (int r = 0; r < W; ++r) {
        if (only >= 0 && r != only) continue;
        RankResult rr = run_rank(seed, dim, n, nq, W, r);
        for (int q = 0; q < nq; ++q) {
            if (mode == "queryx") {
                unsigned b; memcpy(&b, &rr.dist[q], 4);
                printf("%d %d %d %08x\n", q, r, rr.ids[q], b);
            }
            if (r == 0 || only >= 0 || rr.dist[q] < best[q]) best[q] = rr.dist[q];
        }
    }
    if (mode == "lines")
        for (int q = 0; q < nq; ++q) Utility::print_result_line(n + q, best[q]);
    fflush(stdout);
    return 0;
}
