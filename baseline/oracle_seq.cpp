// Oracle B: the reference's own translation units with `main`
// renamed, driven by a tiny command dispatcher.  Nothing algorithmic lives in
// this file — build_tree / nearest_neighbor / generate_problem / print_tree are
// the unmodified functions of baseline/_ref/src.  Test + baseline tooling only.
//
//   oracle_seq MODE SRC DIM N NQ [OUT]
//     SRC  = seed:<int>  (reference generator, N+NQ rows)  |  file:<path> (float32 rows)
//     MODE = tree      reference Utility::print_tree text on stdout
//            head      reference Utility::print_head_and_leaves text
//            points    the generated float32 rows (N+NQ) -> OUT
//            inorder   int32 point IDs in in-order position -> OUT
//            lines     the reference's result lines for the NQ queries
//            queryx    "<query row> <nn id> <dist2 bits hex> <dist bits hex>" per query
//            time      one JSON line with phase timings (generate, build, query)
#define main kd_reference_sequential_main
#include "_ref/src/kdtree_sequential.cpp"
#undef main

#include <cstdio>
#include <cstring>
#include <string>
#include <vector>

static void inorder_ids(Node* n, std::vector<int>& out) {
    // explicit stack: reference trees reach depth 28 only, but keep it iterative anyway
    std::vector<Node*> st;
    Node* cur = n;
    while (cur || !st.empty()) {
        while (cur) { st.push_back(cur); cur = cur->left; }
        cur = st.back(); st.pop_back();
        out.push_back(cur->point->ID);
        cur = cur->right;
    }
}

int main(int argc, char** argv) {
    if (argc < 6) {
        fprintf(stderr, "usage: %s MODE seed:<s>|file:<p> DIM N NQ [OUT]\n", argv[0]);
        return 2;
    }
    std::string mode = argv[1], src = argv[2];
    int dim = atoi(argv[3]), n = atoi(argv[4]), nq = atoi(argv[5]);
    const char* out = argc > 6 ? argv[6] : nullptr;
    auto t0 = std::chrono::high_resolution_clock::now();
    float* x = nullptr;
    if (src.rfind("seed:", 0) == 0) {
        x = Utility::generate_problem(atoi(src.c_str() + 5), dim, n + nq);
    } else if (src.rfind("file:", 0) == 0) {
        size_t cnt = (size_t)(n + nq) * dim;
        x = (float*)calloc(cnt, sizeof(float));
        FILE* f = fopen(src.c_str() + 5, "rb");
        if (!f || fread(x, sizeof(float), cnt, f) != cnt) { fprintf(stderr, "short read\n"); return 2; }
        fclose(f);
    } else { fprintf(stderr, "bad SRC\n"); return 2; }
    auto t1 = std::chrono::high_resolution_clock::now();
    if (mode == "points") {
        FILE* f = fopen(out, "wb");
        fwrite(x, sizeof(float), (size_t)(n + nq) * dim, f);
        fclose(f);
        return 0;
    }
    Point** pts = (Point**)calloc(n, sizeof(Point*));
    for (int i = 0; i < n; ++i) pts[i] = new Point(dim, i + 1, x + (size_t)i * dim);
    auto t2 = std::chrono::high_resolution_clock::now();
    Node* tree = build_tree(pts, n);
    auto t3 = std::chrono::high_resolution_clock::now();

    if (mode == "tree") {
        Utility::print_tree(tree);
    } else if (mode == "head") {
        Utility::print_head_and_leaves(tree);
    } else if (mode == "inorder") {
        std::vector<int> ids; ids.reserve(n);
        inorder_ids(tree, ids);
        FILE* f = fopen(out, "wb");
        fwrite(ids.data(), sizeof(int), ids.size(), f);
        fclose(f);
    } else if (mode == "lines" || mode == "queryx" || mode == "time") {
        std::vector<Node*> res(nq);
        auto q0 = std::chrono::high_resolution_clock::now();
        for (int q = 0; q < nq; ++q) {
            Point query(dim, n + q, x + (size_t)(n + q) * dim);
            res[q] = nearest_neighbor(tree, &query);
        }
        auto q1 = std::chrono::high_resolution_clock::now();
        for (int q = 0; q < nq && mode != "time"; ++q) {
            Point query(dim, n + q, x + (size_t)(n + q) * dim);
            if (mode == "lines") {
                Utility::print_result_line(query.ID, query.distance(*res[q]->point));
            } else {
                float d2 = query.distance_squared(*res[q]->point);
                float d = query.distance(*res[q]->point);
                unsigned b2, b1;
                memcpy(&b2, &d2, 4); memcpy(&b1, &d, 4);
                printf("%d %d %08x %08x\n", q, res[q]->point->ID, b2, b1);
            }
        }
        if (mode == "time") {
            auto s = [](auto a, auto b) { return std::chrono::duration<double>(b - a).count(); };
            printf("{\"n\": %d, \"dim\": %d, \"nq\": %d, \"generate_s\": %.6f, \"wrap_s\": %.6f, "
                   "\"build_s\": %.6f, \"query_s\": %.6f}\n",
                   n, dim, nq, s(t0, t1), s(t1, t2), s(t2, t3), s(q0, q1));
        }
    } else { fprintf(stderr, "bad MODE\n"); return 2; }
    fflush(stdout);
    return 0;  // process exit reclaims the tree
}
