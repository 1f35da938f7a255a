// Test driver of include/directed_scm.hpp: the reference's sample flow (src/bin/sample.rs:80-150) on one
// .flag file, printed as one line of JSON for tests/test_cpp_host.py.
//   host_mirror_check selfcheck                      host maths only (no device)
//   host_mirror_check parse file.flag                read_flag_file only (no device): n, edge count, FNV of the edges
//   host_mirror_check run file.flag seed nsteps simple(0|1)
#include <cstdio>
#include <cstring>
#include "directed_scm.hpp"

using namespace directed_scm;

static void print_vec(const char* key, const Counts& v) {
  printf("\"%s\": [", key);
  for (size_t i = 0; i < v.size(); i++) printf("%s%llu", i ? ", " : "", (unsigned long long)v[i]);
  printf("]");
}

int main(int argc, char** argv) {
  try {
    if (argc >= 2 && !strcmp(argv[1], "selfcheck")) {
      printf("{\"factorial\": [%llu, %llu, %llu, %llu, %llu], \"binomial_5_2\": %llu, ", (unsigned long long)factorial(0),
             (unsigned long long)factorial(1), (unsigned long long)factorial(2), (unsigned long long)factorial(3),
             (unsigned long long)factorial(4), (unsigned long long)binomial(5, 2));
      print_vec("oeis", oeis_a058298());
      printf(", ");
      print_vec("relax_de", calc_relax_de({279, 2200, 2743, 2431, 1839, 1054, 379, 64}));
      printf(", \"devices\": %d, \"abi\": %d}\n", fcm_device_count(), fcm_abi_version());
      return 0;
    }
    if (argc >= 3 && !strcmp(argv[1], "parse")) {
      auto ge = read_flag_file(argv[2]);
      uint64_t h = 1469598103934665603ULL;
      for (const Edge& e : ge.second) for (Node x : e) { h ^= x; h *= 1099511628211ULL; }
      printf("{\"n\": %u, \"nedges\": %zu, \"edges_fnv\": %llu}\n", ge.first, ge.second.size(), (unsigned long long)h);
      return 0;
    }
    if (argc >= 3 && !strcmp(argv[1], "cliques")) {          // sizes of cliques_by_order + an FNV hash of the clique list
      auto ge = read_flag_file(argv[2]);
      const auto by = cliques_by_order(ge.first, ge.second);
      uint64_t h = 1469598103934665603ULL;
      for (const auto& c : compute_maximal_cliques(ge.first, ge.second)) { for (Node x : c) { h ^= x; h *= 1099511628211ULL; } h ^= 0xffu; h *= 1099511628211ULL; }
      printf("{\"by_order\": [");
      for (size_t k = 0; k < by.size(); k++) printf("%s%zu", k ? ", " : "", by[k].size());
      printf("], \"fnv\": %llu}\n", (unsigned long long)h);
      return 0;
    }
    if (argc < 6 || strcmp(argv[1], "run")) { fprintf(stderr, "usage: %s selfcheck | run file.flag seed nsteps simple\n", argv[0]); return 2; }
    auto ge = read_flag_file(argv[2]);
    const uint64_t seed = strtoull(argv[3], 0, 10), nsteps = strtoull(argv[4], 0, 10);
    const bool simple = atoi(argv[5]) != 0;
    Context ctx(0);
    MCMCSampler s = initialize_new_sampler(ctx, ge.first, ge.second, {seed}, 0.01, simple);
    s.step(nsteps);
    ChainView v = s.state(0);
    uint64_t h = 1469598103934665603ULL;                  // FNV-1a over the edge container, in order
    for (const Edge& e : v.edges) for (Node x : e) { h ^= x; h *= 1099511628211ULL; }
    printf("{\"n\": %u, \"nedges\": %zu, \"sample_distance\": %llu, ", ge.first, ge.second.size(), (unsigned long long)s.sample_distance);
    print_vec("bounds_min", s.bounds.flag_count_min); printf(", ");
    print_vec("bounds_max", s.bounds.flag_count_max); printf(", ");
    print_vec("flag_count", v.flag_count);
    printf(", \"sampled\": %llu, \"accepted\": %llu, \"edges_fnv\": %llu, \"bounds_ok\": %d}\n", (unsigned long long)v.sampled,
           (unsigned long long)v.accepted, (unsigned long long)h, (int)s.bounds.check(v.flag_count));
    return 0;
  } catch (const Error& e) {
    fprintf(stderr, "error %d: %s\n", e.code, e.what());
    return 1;
  }
}
