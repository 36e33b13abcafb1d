// cli.cpp -- `founderblockgraph`: the reference's command line in front of
// libfbg_cuda.so.  Same flags, defaults, validation order, messages and exit codes
// as founderblockgraph_cmdline.ggo:23-36 and main() (fbg.cpp:3298-3521); the work
// between "read the MSA" and "write the xGFA" is done on the GPU through the C ABI
// (include/fbg_cuda.h).  There is no CPU fallback.
//
// Deliberate differences (DESIGN.md "Out of scope / deviations"):
//  * no <input>.plain / <input>.plain<gl>.cst side files (fbg.cpp:363-365,430);
//  * the "index requires X MiB" line reports the flat device arrays, not an sdsl CST;
//  * --threads is validated like the reference (fbg.cpp:3396-3399) and otherwise unused;
//  * non-elastic mode: the reference prints its statistics and then crashes in
//    make_efg.  This tool prints the same statistics and stops with exit 1,
//    unless the extension flag --nonelastic-gfa is given, in which case it writes the
//    xGFA of that segmentation with the elastic writer.
//  * --heuristic-subset=ROWNUM (hidden; fbg.cpp:3400-3433, 3467-3500) needs --elastic --gfa: f(x) from row chunks
//    (one chunk's index in GPU memory at a time), one DP, then the repair loop; same xGFA bytes as the reference
//    for m <= 100 rows (beyond that the reference's transpose worker corrupts its own input, fbg.cpp:330-337).
//    With -p the reference runs into an assertion of its own while writing (fbg.cpp:1421): so does this tool
//    (message, exit 134, no file).  Always on one GPU.
//
// Several GPUs (the reference's counterpart is --threads, fbg.cpp:3390-3399): FBG_GPUS=N (or "all") builds ONE
// graph on N GPUs of the box.  After the FASTA is parsed -- and before this process touches CUDA -- the tool
// forks N - 1 ranks (they share the parsed MSA copy-on-write); every rank opens its own GPU, joins the NCCL
// communicator whose id rank 0 hands out over pipes, and makes the same C-ABI calls (sharded index, f and DP,
// include/fbg_cuda.h "one MSA across the GPUs of a box"); rank 0 alone prints and writes the file.
#include "../../../include/fbg_cuda.h"
#include <getopt.h>
#include <sys/types.h>
#include <sys/wait.h>
#include <unistd.h>
#include <csignal>
#include <chrono>
#include <thread>
#include <utility>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <string>
#include <vector>

namespace {

const char *USAGE =
	"Usage: founderblockgraph --input=MSA.fasta --output={MSA.index|efg.xgfa} [--gfa]\n"
	"[--elastic] [--gap-limit=GAPLIMIT] [--threads=THREADNUM]\n"
	"[--graphviz-output=efg.dot] [--output-paths] [--ignore-chars=\"ALPHABET\"]";
const char *PURPOSE = "Constructs a semi-repeat-free (Elastic) Founder Graph";
const char *DESCRIPTION =
	"Input is MSA given in fasta format. In standard mode (without --elastic), rows\n"
	"with runs of gaps \xE2\x80\x98-\xE2\x80\x99 or N\xE2\x80\x99s \xE2\x89\xA5 GAPLIMIT will be filtered out.";
const char *HELP[] = {
	"  -h, --help                    Print help and exit",
	"      --full-help               Print help, including hidden options, and exit",
	"  -V, --version                 Print version and exit",
	"      --input=filename          MSA input path",
	"      --output=filename         Index/EFG output path",
	"      --gap-limit=GAPLIMIT      Gap limit (suppressed by --elastic)\n                                  (default=`1')",
	"      --graphviz-output=filename\n                                Graphviz output path",
	"      --memory-chart-output=filename\n                                Memory chart output path",
	"  -e, --elastic                 Min-max-length semi-repeat-free segmentation\n                                  (default=off)",
	"      --gfa                     Saves output in xGFA format  (default=off)",
	"  -p, --output-paths            Print the original sequences as paths of the\n                                  xGFA graph (requires --gfa)  (default=off)",
	"      --ignore-chars=STRING     Ignore these characters for the indexability\n                                  property/pattern matching",
	"  -t, --threads=THREADNUM       Max # threads  (default=`-1')",
	"      --heuristic-subset=ROWNUM To save memory, compute the optimal\n                                  segmentation in chunks of ROWNUM MSA rows,\n                                  then fix the resulting graph iteratively,\n                                  sacrificing optimality  (default=`-1')",
	"      --disable-elastic-tricks  Disable the tricks considering the start and\n                                  end of sequences as unique  (default=off)",
};

void print_help(bool full) {
	printf("%s\n%s\n\n%s\n\n", USAGE, PURPOSE, DESCRIPTION);
	for (int i = 0; i < (full ? 15 : 13); i++) printf("%s\n", HELP[i]);
}

struct Args {
	std::string input, output, graphviz, ignore;
	bool input_given = false, output_given = false, graphviz_given = false, ignore_given = false;
	long gap_limit = 1, threads = -1, heuristic_subset = -1;
	bool elastic = false, gfa = false, paths = false, tricks_off = false, nonelastic_gfa = false;
};

bool parse_long(const char *prog, const char *val, long *out) {
	char *stop = nullptr;
	*out = strtol(val, &stop, 0);
	if (!(stop && *stop == '\0')) { fprintf(stderr, "%s: invalid numeric value: %s\n", prog, val); return false; }
	return true;
}

// returns 0 ok, 1 failure (exit 1), 2 exit 0 (help/version)
int parse_args(int argc, char **argv, Args &a) {
	static struct option long_options[] = {
		{"help", 0, nullptr, 'h'}, {"full-help", 0, nullptr, 0}, {"version", 0, nullptr, 'V'},
		{"input", 1, nullptr, 0}, {"output", 1, nullptr, 0}, {"gap-limit", 1, nullptr, 0},
		{"graphviz-output", 1, nullptr, 0}, {"memory-chart-output", 1, nullptr, 0},
		{"elastic", 0, nullptr, 'e'}, {"gfa", 0, nullptr, 0}, {"output-paths", 0, nullptr, 'p'},
		{"ignore-chars", 1, nullptr, 0}, {"threads", 1, nullptr, 't'}, {"heuristic-subset", 1, nullptr, 0},
		{"disable-elastic-tricks", 0, nullptr, 0}, {"nonelastic-gfa", 0, nullptr, 0}, {nullptr, 0, nullptr, 0}};
	const char *prog = argv[0];
	int given[16] = {0};
	auto once = [&](int idx, const char *name, char shortc) -> bool {
		if (given[idx]++) {
			if (shortc != '-') fprintf(stderr, "%s: `--%s' (`-%c') option given more than once\n", prog, name, shortc);
			else fprintf(stderr, "%s: `--%s' option given more than once\n", prog, name);
			return false;
		}
		return true;
	};
	optarg = nullptr; optind = 0; opterr = 1; optopt = '?';
	for (;;) {
		int idx = 0;
		int c = getopt_long(argc, argv, "hVept:", long_options, &idx);
		if (c == -1) break;
		switch (c) {
		case 'h': print_help(false); return 2;
		case 'V': printf("founderblockgraphs 0.5\n"); return 2;
		case 'e': if (!once(8, "elastic", 'e')) return 1; a.elastic = !a.elastic; break;
		case 'p': if (!once(10, "output-paths", 'p')) return 1; a.paths = !a.paths; break;
		case 't': if (!once(12, "threads", 't')) return 1; if (!parse_long(prog, optarg, &a.threads)) return 1; break;
		case 0: {
			const char *nm = long_options[idx].name;
			if (!strcmp(nm, "full-help")) { print_help(true); return 2; }
			else if (!strcmp(nm, "input")) { if (!once(3, nm, '-')) return 1; a.input = optarg; a.input_given = true; }
			else if (!strcmp(nm, "output")) { if (!once(4, nm, '-')) return 1; a.output = optarg; a.output_given = true; }
			else if (!strcmp(nm, "gap-limit")) { if (!once(5, nm, '-')) return 1; if (!parse_long(prog, optarg, &a.gap_limit)) return 1; }
			else if (!strcmp(nm, "graphviz-output")) { if (!once(6, nm, '-')) return 1; a.graphviz = optarg; a.graphviz_given = true; }
			else if (!strcmp(nm, "memory-chart-output")) { if (!once(7, nm, '-')) return 1; }
			else if (!strcmp(nm, "gfa")) { if (!once(9, nm, '-')) return 1; a.gfa = !a.gfa; }
			else if (!strcmp(nm, "ignore-chars")) { if (!once(11, nm, '-')) return 1; a.ignore = optarg; a.ignore_given = true; }
			else if (!strcmp(nm, "heuristic-subset")) { if (!once(13, nm, '-')) return 1; if (!parse_long(prog, optarg, &a.heuristic_subset)) return 1; }
			else if (!strcmp(nm, "disable-elastic-tricks")) { if (!once(14, nm, '-')) return 1; a.tricks_off = !a.tricks_off; }
			else if (!strcmp(nm, "nonelastic-gfa")) { if (!once(15, nm, '-')) return 1; a.nonelastic_gfa = true; }
			break;
		}
		case '?': return 1; // getopt_long already printed the message
		default: abort();
		}
	}
	int err = 0;
	if (!a.input_given) { fprintf(stderr, "%s: '--input' option required\n", prog); err = 1; }
	if (!a.output_given) { fprintf(stderr, "%s: '--output' option required\n", prog); err = 1; }
	return err;
}

void write_graphviz_empty(const char *path) {
	// elastic mode never fills node_labels/edges (fbg.cpp:3503,3511-3515): empty digraph
	FILE *fp = fopen(path, "wb");
	if (!fp) return;
	fputs("digraph founder_block_graph {\nrankdir=\"LR\"\n}\n", fp);
	fclose(fp);
}

int g_rank = 0; // > 0: a forked rank of a multi-GPU run (silent: rank 0 speaks for all)
volatile sig_atomic_t g_kids_done = 0;

// rank 0: a forked rank that dies would leave the others waiting in a collective for ever
void on_sigchld(int) {
	int st = 0;
	while (waitpid(-1, &st, WNOHANG) > 0) {
		if (!WIFEXITED(st) || WEXITSTATUS(st) != 0) {
			static const char msg[] = "ERROR: a GPU rank of this run failed\n";
			if (write(2, msg, sizeof msg - 1) < 0) {}
			_exit(EXIT_FAILURE);
		}
		g_kids_done = g_kids_done + 1;
	}
}

int fail_cuda(fbg_ctx *ctx, const char *what) {
	fprintf(stderr, "ERROR: %s%s: %s\n", g_rank ? "(GPU rank > 0) " : "", what, fbg_last_error(ctx));
	return EXIT_FAILURE;
}

// GPUs visible to CUDA, asked in a short-lived child so that this process stays fork-safe (no CUDA state yet)
int visible_gpus() {
	int fd[2];
	if (pipe(fd) != 0) return 1;
	pid_t pid = fork();
	if (pid == 0) {
		close(fd[0]);
		int n = fbg_device_count();
		if (write(fd[1], &n, sizeof n) != (ssize_t) sizeof n) _exit(1);
		_exit(0);
	}
	close(fd[1]);
	int n = 1;
	if (pid < 0 || read(fd[0], &n, sizeof n) != (ssize_t) sizeof n) n = 1;
	close(fd[0]);
	if (pid > 0) waitpid(pid, nullptr, 0);
	return n < 1 ? 1 : n;
}

// FBG_GPUS: unset / 1 = one GPU, N = that many, "all" = every visible one
int wanted_gpus() {
	const char *e = getenv("FBG_GPUS");
	if (!e || !*e) return 1;
	if (!strcmp(e, "all")) return visible_gpus();
	long v = strtol(e, nullptr, 10);
	return v < 1 ? 1 : (v > 64 ? 64 : (int) v);
}

} // namespace

int main(int argc, char **argv) {
	Args a;
	int pr = parse_args(argc, argv, a);
	if (pr == 2) return EXIT_SUCCESS;
	if (pr) return EXIT_FAILURE;

	// fbg.cpp:3308-3334, same order
	if (a.gap_limit < 0) { fprintf(stderr, "Gap limit needs to be non-negative.\n"); return EXIT_FAILURE; }
	if (!a.elastic && a.paths) { fprintf(stderr, "Output of original sequences as paths without option --elastic is not implemented!\n"); return EXIT_FAILURE; }
	if (!a.nonelastic_gfa && ((!a.elastic && a.gfa) || (a.elastic && !a.gfa))) {
		fprintf(stderr, "--elastic and --gfa options are currently only supported when both are used!\n");
		return EXIT_FAILURE;
	}
	if (a.heuristic_subset < -1 || a.heuristic_subset == 0) { fprintf(stderr, "wrong value for --heuristic-subset!\n"); return EXIT_FAILURE; }
	const bool hs = a.heuristic_subset != -1;
	if (hs && !(a.elastic && a.gfa)) { fprintf(stderr, "--heuristic-subset is only supported together with --elastic --gfa.\n"); return EXIT_FAILURE; }

	auto start = std::chrono::high_resolution_clock::now();
	// FBG_TIMING=1: wall-clock per stage on stderr at the end (parse and write are host work)
	const bool timing = getenv("FBG_TIMING") != nullptr;
	std::vector<std::pair<const char *, double>> laps;
	auto last = start;
	auto lap = [&](const char *what) {
		auto now = std::chrono::high_resolution_clock::now();
		laps.emplace_back(what, std::chrono::duration<double, std::milli>(now - last).count());
		last = now;
	};
	// the CUDA context takes a few hundred milliseconds to come up: create it while the FASTA is parsed
	// (one GPU; with several, the ranks are forked after the parse and no CUDA call may come before that)
	const int world = hs ? 1 : wanted_gpus();
	fbg_ctx *ctx = nullptr;
	int ctx_rc = FBG_OK;
	std::thread ctx_thread([&]() { if (world == 1) ctx_rc = fbg_ctx_create(&ctx, -1); });
	// read_input (fbg.cpp:151-201): host/fasta_reader.cpp, mapped file, several threads
	fbg_msa msa;
	bool opened = fbg_read_fasta(a.input.c_str(), (uint64_t) a.gap_limit, a.elastic, a.paths, 0, &msa) == FBG_OK;
	if (opened && hs && msa.messages && msa.messages[0]) { // parse_input takes every row or none (fbg.cpp:230-234, 250-254)
		ctx_thread.join();
		fprintf(stderr, "MSA rows have mismatching size!\n");
		return EXIT_FAILURE;
	}
	if (opened && msa.messages) fputs(msa.messages, stderr);
	if (!opened || msa.m == 0 || msa.n == 0) { ctx_thread.join(); fprintf(stderr, "Unable to read sequences from the input\n."); return EXIT_FAILURE; }
	fprintf(stderr, "Input MSA[1..%llu,1..%llu]\n", (unsigned long long) msa.m, (unsigned long long) msa.n);

	lap("parse");
	ctx_thread.join();
	std::vector<pid_t> kids;
	if (world > 1) {
		// fork the other ranks (this process is rank 0), then hand out the NCCL id
		std::vector<int> wfd;
		int my_rfd = -1;
		fflush(stdout); fflush(stderr);
		signal(SIGCHLD, on_sigchld);
		for (int r = 1; r < world; r++) {
			int fd[2];
			if (pipe(fd) != 0) { fprintf(stderr, "ERROR: cannot create a pipe for GPU rank %d\n", r); return EXIT_FAILURE; }
			pid_t pid = fork();
			if (pid < 0) { fprintf(stderr, "ERROR: cannot fork GPU rank %d\n", r); return EXIT_FAILURE; }
			if (pid == 0) {
				g_rank = r; my_rfd = fd[0]; close(fd[1]);
				for (int w : wfd) close(w);
				kids.clear();
				signal(SIGCHLD, SIG_DFL);
				if (!getenv("FBG_DEBUG") && !freopen("/dev/null", "w", stderr)) _exit(1); // rank 0 speaks for all
				break;
			}
			close(fd[0]); wfd.push_back(fd[1]); kids.push_back(pid);
		}
		ctx_rc = fbg_ctx_create(&ctx, g_rank);
		if (ctx_rc != FBG_OK) { int rc = fail_cuda(nullptr, "cannot initialise the GPU"); if (g_rank) _exit(rc); return rc; }
		uint8_t id[128];
		if (g_rank == 0) {
			if (fbg_dist_unique_id(id) != FBG_OK) return fail_cuda(nullptr, "NCCL");
			for (int w : wfd) { if (write(w, id, sizeof id) != (ssize_t) sizeof id) { fprintf(stderr, "ERROR: cannot reach a GPU rank\n"); return EXIT_FAILURE; } close(w); }
		} else {
			size_t got = 0;
			while (got < sizeof id) { ssize_t k = read(my_rfd, id + got, sizeof id - got); if (k <= 0) _exit(1); got += (size_t) k; }
			close(my_rfd);
		}
		if (fbg_dist_init(ctx, g_rank, world, id) != FBG_OK) { int rc = fail_cuda(ctx, "joining the GPUs of this run"); if (g_rank) _exit(rc); return rc; }
		if (g_rank == 0) fprintf(stderr, "Building on %d GPUs.\n", world);
	}
	if (ctx_rc != FBG_OK) return fail_cuda(nullptr, "cannot initialise the GPU");
	lap("GPU context (rest after the parse)");
	if (fbg_set_options(ctx, a.elastic, a.gap_limit, (const uint8_t *) a.ignore.data(), a.ignore.size(), a.tricks_off) != FBG_OK) return fail_cuda(ctx, "options");
	fbg_stats st;
	if (hs) {
		// fbg.cpp:3400-3433: f(x) chunk by chunk, the maximum of the chunks' values
		const uint64_t R = (uint64_t) a.heuristic_subset, nch = (msa.m + R - 1) / R;
		std::vector<uint64_t> cg(nch), ci(nch);
		int rc = fbg_heuristic_f(ctx, msa.rows, msa.m, msa.n, R, cg.data(), ci.data());
		if (rc != FBG_OK) return fail_cuda(ctx, "index construction (row chunks)");
		for (uint64_t k = 0; k < nch; k++) {
			fprintf(stderr, "Reading MSA[%llu..%llu]...\n", (unsigned long long) (k * R), (unsigned long long) (k * R + R - 1));
			fprintf(stderr, "MSA contains %llu gaps.\n", (unsigned long long) cg[k]);
			fprintf(stderr, "MSA contains %llu characters to ignore for the semi-repeat-free property.\n", (unsigned long long) ci[k]);
		}
		fbg_get_stats(ctx, &st);
		lap("row-chunk indexes + f");
	} else {
	if (fbg_upload_msa(ctx, msa.rows, msa.m, msa.n) != FBG_OK) return fail_cuda(ctx, "MSA upload");
	fprintf(stderr, "Building the MSA index (suffix array, LCP) on the GPU.\n");
	// index + f(x) in one call: the library then builds only as much of the index as f(x) can
	// depend on (on gapless input far less than the full SA / ISA / LCP an explicit fbg_build_index gives)
	if (fbg_compute_f(ctx) != FBG_OK) return fail_cuda(ctx, "index construction");
	fbg_get_stats(ctx, &st);
	lap("upload + index + f");
	// the reference reports size_in_mega_bytes(cst) here (fbg.cpp:3380); this build reports the device memory its index
	// stage took (persistent arrays + scratch peak), which is 0-array "virtual" on gapless input
	fprintf(stderr, "MSA index construction complete, index requires %g MiB.\n", (double) st.arena_peak_bytes / (1024.0 * 1024.0));
	// fbg.cpp:3396-3399: checked only now, after the index has been built
	if (a.elastic && !(a.threads == -1 || a.threads > 0)) { fprintf(stderr, "Invalid number of threads.\n"); return EXIT_FAILURE; }
	}

	uint64_t nb = 0, score = 0;
	fbg_graph g;
	memset(&g, 0, sizeof g);
	std::vector<uint64_t> bnd;
	if (a.elastic) {
		if (!hs) {
			fprintf(stderr, "MSA contains %llu gaps.\n", (unsigned long long) st.gaps);
			fprintf(stderr, "MSA contains %llu characters to ignore for the semi-repeat-free property.\n", (unsigned long long) st.to_ignore);
		}
		int rc = fbg_segment(ctx, &nb, &score);
		if (rc == FBG_ERR_NOSEG) { fprintf(stderr, "No valid segmentation found!\n"); return EXIT_FAILURE; }
		if (rc != FBG_OK) return fail_cuda(ctx, "segmentation");
		fprintf(stderr, "Computing optimal segmentation...done (optimal segment length = %llu).\n", (unsigned long long) score);
	} else {
		int rc = fbg_segment(ctx, &nb, &score);
		if (rc != FBG_OK && rc != FBG_ERR_NOSEG) return fail_cuda(ctx, "segmentation");
		fprintf(stderr, "Optimal score: %llu\n", (unsigned long long) score);
		if (rc == FBG_ERR_NOSEG) {
			fprintf(stderr, a.gap_limit == 1 ? "No proper segmentation exists.\n" : "No valid segmentation found!\n");
			return EXIT_FAILURE;
		}
		fprintf(stderr, "Number of segments: %llu\n", (unsigned long long) nb);
	}
	// the collective stages are over: graph and file are rank 0's (K4 / K5 use no collective); the other ranks leave,
	// rank 0 waits until they have (the watchdog above ends the run if one of them failed)
	if (g_rank) _exit(EXIT_SUCCESS);
	while (world > 1 && g_kids_done < (sig_atomic_t) kids.size()) usleep(200);
	if (hs) {
		// fbg.cpp:3467-3500: merge blocks until every node passes efg_validate
		int iterations = 0;
		for (;;) {
			uint64_t removed = 0, left = 0;
			fbg_graph gc;
			iterations++;
			if (fbg_validate_blocks(ctx, &removed, &left, &gc) != FBG_OK) return fail_cuda(ctx, "graph validation");
			fprintf(stderr, "#nodes=%llu\ntotal length of node labels=%llu\n#edges=%llu\n", (unsigned long long) gc.n_nodes, (unsigned long long) gc.label_total, (unsigned long long) gc.n_edges);
			fprintf(stderr, "There are %llu blocks to remove\n", (unsigned long long) removed);
			nb = left;
			if (removed == 0) break;
		}
		fprintf(stderr, "Graph fixed in %diterations\xE2\x80\xA6\n", iterations - 1);
		lap("repair loop");
	}
	bnd.resize(nb);
	if (fbg_get_boundaries(ctx, bnd.data()) != FBG_OK) return fail_cuda(ctx, "boundaries");
	lap("segmentation");
	// Graph and xGFA bytes: formatted on the device and streamed to the file (K5, default), or --
	// FBG_HOST_WRITER=1 -- graph arrays copied out and formatted by the host writer.  Same bytes.
	const bool host_writer = getenv("FBG_HOST_WRITER") != nullptr;
	const bool ids_bad = a.paths && msa.n_ids != msa.m; // the reference asserts at the end of output_efg (fbg.cpp:1292)
	const char *const *ids = msa.ids;
	uint64_t xgfa_bytes = 0;
	if (host_writer) {
		if (fbg_build_graph(ctx, a.paths ? 1 : 0, &g) != FBG_OK) return fail_cuda(ctx, "graph construction");
	} else if (!ids_bad) {
		if (fbg_format_xgfa(ctx, a.paths ? 1 : 0, a.paths ? ids : nullptr, &g, &xgfa_bytes) != FBG_OK) return fail_cuda(ctx, "graph construction");
	}
	lap(host_writer ? "graph" : "graph + xGFA bytes");

	if (!a.elastic) {
		fprintf(stderr, "#nodes=%llu\n", (unsigned long long) g.n_nodes);
		fprintf(stderr, "total length of node labels=%llu\n", (unsigned long long) g.label_total);
		fprintf(stderr, "#founders=%llu\n", (unsigned long long) g.max_height);
		fprintf(stderr, "#edges=%llu\n", (unsigned long long) g.n_edges);
		if (!a.nonelastic_gfa) {
			fprintf(stderr, "Writing the index to disk\xE2\x80\xA6\n");
			fprintf(stderr, "ERROR: the reference crashes at this point (make_efg on an empty segmentation, fbg.cpp:3449) and writes nothing; "
			                "rerun with --nonelastic-gfa to write this segmentation as xGFA.\n");
			return EXIT_FAILURE;
		}
	}
	fprintf(stderr, "Writing the xGFA to disk\xE2\x80\xA6\n");
	if (hs && a.paths) {
		fprintf(stderr, "founderblockgraph: boundaries[j] > previndex - 1 failed (the reference's output_efg_external cannot write paths, fbg.cpp:1421)\n");
		return 134;
	}
	if (ids_bad) {
		fprintf(stderr, "founderblockgraph: identifiers.size() == paths.size() failed (rows were skipped while reading; fbg.cpp:1292)\n");
		return 134;
	}
	const int wrc = host_writer ? fbg_write_xgfa(a.output.c_str(), msa.rows, msa.m, msa.n, bnd.data(), &g, a.paths ? 1 : 0, a.paths ? ids : nullptr)
	                            : fbg_save_xgfa(ctx, a.output.c_str());
	if (wrc != FBG_OK) {
		fprintf(stderr, "ERROR: cannot write %s\n", a.output.c_str());
		return EXIT_FAILURE;
	}
	lap("xGFA write");
	fbg_graph_free(&g);
	fbg_msa_free(&msa);
	fbg_ctx_destroy(ctx);
	lap("teardown");
	if (timing) {
		double tot = 0;
		fprintf(stderr, "[fbg] stages (ms):");
		for (auto &l : laps) { fprintf(stderr, " %s %.1f,", l.first, l.second); tot += l.second; }
		fprintf(stderr, " total %.1f\n", tot);
	}

	auto end = std::chrono::high_resolution_clock::now();
	auto secs = std::chrono::duration_cast<std::chrono::seconds>(end - start).count();
	if (a.graphviz_given) {
		fprintf(stderr, "Writing the Graphviz file\xE2\x80\xA6\n");
		write_graphviz_empty(a.graphviz.c_str());
	}
	fprintf(stderr, "Time taken: %lld seconds\n", (long long) secs);
	return EXIT_SUCCESS;
}
