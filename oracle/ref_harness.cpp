// oracle/ref_harness.cpp -- TEST INFRASTRUCTURE ONLY. Never linked into, called
// from, or shipped with the product (founderblockgraphs_b200/). Only tests/,
// __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs
// may execute the binary built from this file.
//
// What it is: a driver around the UNMODIFIED reference translation unit
// (/root/reference/founderblockgraph.cpp, included via the sed-generated
// oracle/_ref/fbg_instrumented.cpp which only adds two dump hooks, see
// instrument.sed).  All reference functions live in an anonymous namespace
// (fbg.cpp:38), so same-TU inclusion is the only way to call them
// (SURVEY.md 8(c) O4).  The reference's main() is renamed and never called.
//
// Why: at the reference commit the non-elastic path crashes after segment()
//  and the elastic path exposes neither f nor the DP tables.  This
// harness calls read_input / load_cst / segment / segment2elasticValid /
// segment_elastic_minmaxlength{,_multithread} / output_efg exactly as main()
// does (fbg.cpp:3349-3503) and prints the intermediate vectors on stdout:
//
//   m <rows>            n <cols>
//   f <n values>                         (elastic)
//   mml <n+1 values>  backtrack <n+1>    (elastic)
//   boundaries <B values>                (both; elastic: last element is n)
//   v <n> / s <n> / prev <n>             (non-elastic)
//   status <exit status the reference main() would return before writing>
//   time_index_s / time_segment_s / time_write_s   (wall clock, for cpu_baseline)
//
// The reference's own stderr lines pass through untouched.

#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <chrono>
#include <iostream>
#include <string>
#include <vector>

static bool g_dump = true;

template <class V>
static void dump_vec(const char *name, const V &v)
{
	if (!g_dump) return;
	std::printf("%s", name);
	for (auto const x : v) std::printf(" %llu", (unsigned long long) x);
	std::printf("\n");
}

#define FBG_ORACLE_DUMP_NONELASTIC(v, s, prev, boundaries) \
	do { dump_vec("v", v); dump_vec("s", s); dump_vec("prev", prev); dump_vec("boundaries", boundaries); } while (0)
#define FBG_ORACLE_DUMP_ELASTIC(mml, backtrack) \
	do { dump_vec("mml", mml); dump_vec("backtrack", backtrack); } while (0)

#define main fbg_reference_main
#include "fbg_instrumented.cpp"
#undef main

static double now_s()
{
	return std::chrono::duration<double>(std::chrono::steady_clock::now().time_since_epoch()).count();
}

int main(int argc, char **argv)
{
	std::string mode, input, output, ignorechars;
	long gap_limit = 1, threads = -1;
	bool tricks_off = false, paths = false;
	for (int a = 1; a < argc; a++) {
		std::string s(argv[a]);
		auto next = [&]() -> std::string { if (a + 1 >= argc) { std::fprintf(stderr, "missing value for %s\n", s.c_str()); std::exit(2); } return std::string(argv[++a]); };
		if (s == "--mode") mode = next();
		else if (s == "--input") input = next();
		else if (s == "--output") output = next();
		else if (s == "--ignore-chars") ignorechars = next();
		else if (s == "--gap-limit") gap_limit = std::atol(next().c_str());
		else if (s == "--threads") threads = std::atol(next().c_str());
		else if (s == "--disable-elastic-tricks") tricks_off = true;
		else if (s == "--output-paths") paths = true;
		else if (s == "--no-dump") g_dump = false;
		else { std::fprintf(stderr, "unknown argument %s\n", s.c_str()); return 2; }
	}
	if (input.empty() || (mode != "elastic" && mode != "nonelastic")) {
		std::fprintf(stderr, "usage: ref_harness --mode elastic|nonelastic --input F [--gap-limit G] [--ignore-chars S] [--disable-elastic-tricks] [--threads T] [--output X] [--output-paths] [--no-dump]\n");
		return 2;
	}
	const bool elastic = (mode == "elastic");

	// fbg.cpp:3345-3359
	std::vector<std::string> MSA, identifiers;
	read_input(input.c_str(), (std::size_t) gap_limit, elastic, MSA, paths, identifiers);
	if (MSA.empty()) {
		std::cerr << "Unable to read sequences from the input\n.";
		std::printf("status 1\n");
		return 0;
	}
	size_type const m = MSA.size(), n = MSA[0].size();
	std::printf("m %llu\nn %llu\n", (unsigned long long) m, (unsigned long long) n);

	// fbg.cpp:3369-3381
	double t0 = now_s();
	cst_type cst;
	if (!load_cst(input.c_str(), MSA, cst, gap_limit)) { std::printf("status 1\n"); return 0; }
	double t1 = now_s();
	std::printf("time_index_s %.6f\n", t1 - t0);

	std::vector<std::string> node_labels;
	adjacency_list edges;
	std::vector<size_type> block_indices;
	std::vector<size_type> f(n, 0);
	int status = EXIT_SUCCESS;
	if (elastic) {
		// fbg.cpp:3392-3399. exit(1) inside (tricks off, f[0]==n) terminates the harness too.
		if (threads == -1)
			segment_elastic_minmaxlength(MSA, cst, ignorechars, block_indices, tricks_off, f);
		else
			segment_elastic_minmaxlength_multithread(MSA, cst, ignorechars, block_indices, threads, tricks_off, f);
		dump_vec("f", f);
		dump_vec("boundaries", block_indices);
	} else if (gap_limit == 1) {
		status = segment(MSA, cst, node_labels, edges);          // fbg.cpp:3437
	} else {
		status = segment2elasticValid(MSA, cst, node_labels, edges); // fbg.cpp:3439
	}
	double t2 = now_s();
	std::printf("time_segment_s %.6f\n", t2 - t1);
	std::printf("status %d\n", status == EXIT_SUCCESS ? 0 : 1);

	if (elastic && !output.empty()) {
		output_efg(block_indices, MSA, paths, identifiers, output.c_str()); // fbg.cpp:3503
		std::printf("time_write_s %.6f\n", now_s() - t2);
	}
	std::fflush(stdout);
	return 0;
}
