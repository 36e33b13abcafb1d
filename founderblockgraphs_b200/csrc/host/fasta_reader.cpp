// fasta_reader.cpp -- read_input / check_sequence_length / check_gaps of the reference
// (fbg.cpp:103-201) for large inputs: the file is mapped, the header lines
// are found by all threads at once (memchr over disjoint byte ranges), and the records are
// then measured, filtered and copied into the row matrix in parallel.  Same acceptance rules,
// same messages in the same order:
//   * the first line is a header whatever it holds (fbg.cpp:160); later lines are headers iff
//     their first byte is '>' (fbg.cpp:165); every other line is sequence data, bytes verbatim
//     ('\r' included), line breaks removed;
//   * a record is kept iff its length equals the FIRST record's (fbg.cpp:176,196) and, without
//     --elastic, its longest run of '-' is below gap_limit (gap_limit 0 keeps all; fbg.cpp:103-133);
//   * with output_paths the header list holds EVERY header, kept or not (fbg.cpp:161,169).
#include "../../../include/fbg_cuda.h"
#include <atomic>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <fcntl.h>
#include <memory>
#include <string>
#include <sys/mman.h>
#include <sys/stat.h>
#include <thread>
#include <unistd.h>
#include <vector>

namespace {

struct MsaHolder {
	std::unique_ptr<uint8_t[]> rows;
	std::vector<std::string> ids;
	std::vector<const char *> id_ptrs;
	std::string messages;
};

struct Record {
	size_t head, head_end; // header line [head, head_end)
	size_t seq, seq_end;   // sequence lines [seq, seq_end), line breaks included
	size_t len = 0;        // sequence bytes without line breaks
	size_t gaprun = 0;     // longest run of '-'
	bool keep = false;
	size_t row = 0;
};

template <class F> void parallel_for(size_t count, unsigned threads, size_t grain, F f) {
	if (threads <= 1 || count <= grain) { for (size_t i = 0; i < count; i++) f(i); return; }
	std::atomic<size_t> next(0);
	auto work = [&]() {
		for (;;) {
			size_t b = next.fetch_add(grain);
			if (b >= count) break;
			size_t e = b + grain < count ? b + grain : count;
			for (size_t i = b; i < e; i++) f(i);
		}
	};
	std::vector<std::thread> th;
	for (unsigned t = 1; t < threads; t++) th.emplace_back(work);
	work();
	for (auto &x : th) x.join();
}

// length without line breaks; longest '-' run (runs continue across line breaks, as in the
// reference, which concatenates the lines first)
void measure(const char *base, Record &r, bool want_gaps) {
	const char *p = base + r.seq, *end = base + r.seq_end;
	size_t breaks = 0;
	for (const char *q = p; q < end;) {
		const char *nl = (const char *) memchr(q, '\n', (size_t) (end - q));
		if (!nl) break;
		breaks++; q = nl + 1;
	}
	r.len = (size_t) (end - p) - breaks;
	r.gaprun = 0;
	if (want_gaps && p < end && memchr(p, '-', (size_t) (end - p))) {
		size_t run = 0, best = 0;
		for (const char *q = p; q < end; q++) {
			const char c = *q;
			if (c == '-') run++;
			else if (c != '\n') { if (run > best) best = run; run = 0; }
		}
		r.gaprun = run > best ? run : best;
	}
}

void copy_record(const char *base, const Record &r, uint8_t *dst) {
	const char *q = base + r.seq, *end = base + r.seq_end;
	while (q < end) {
		const char *nl = (const char *) memchr(q, '\n', (size_t) (end - q));
		const size_t k = (size_t) ((nl ? nl : end) - q);
		memcpy(dst, q, k);
		dst += k;
		if (!nl) break;
		q = nl + 1;
	}
}

} // namespace

extern "C" int fbg_read_fasta(const char *path, uint64_t gap_limit, int elastic, int output_paths, int threads, fbg_msa *out) {
	if (!path || !out) return FBG_ERR_ARG;
	memset(out, 0, sizeof *out);
	int fd = open(path, O_RDONLY);
	if (fd < 0) return FBG_ERR_ARG;
	struct stat sb;
	if (fstat(fd, &sb) != 0) { close(fd); return FBG_ERR_ARG; }
	size_t fsize = 0;
	const char *base = nullptr;
	bool mapped = false;
	std::vector<char> piped; // pipes, process substitution: no mapping, read to the end
	if (S_ISREG(sb.st_mode) && sb.st_size > 0) {
		fsize = (size_t) sb.st_size;
		void *mp = mmap(nullptr, fsize, PROT_READ, MAP_PRIVATE | MAP_POPULATE, fd, 0);
		if (mp != MAP_FAILED) { base = static_cast<const char *>(mp); mapped = true; }
	}
	if (!mapped) {
		char buf[1 << 16];
		for (;;) {
			ssize_t got = read(fd, buf, sizeof buf);
			if (got < 0) { close(fd); return FBG_ERR_ARG; }
			if (got == 0) break;
			piped.insert(piped.end(), buf, buf + got);
		}
		base = piped.data(); fsize = piped.size();
	}
	close(fd);

	unsigned T = threads > 0 ? (unsigned) threads : std::thread::hardware_concurrency();
	if (T < 1) T = 1;
	if (threads <= 0) { if (T > 16) T = 16; if (fsize < (4u << 20)) T = 1; }
	if (T > 256) T = 256;

	// 1. header lines: the first line, and every later line whose first byte is '>'
	std::vector<std::vector<size_t>> found(T);
	{
		auto scan = [&](unsigned t) {
			size_t s = (size_t) ((unsigned __int128) fsize * t / T), e = (size_t) ((unsigned __int128) fsize * (t + 1) / T);
			// the lines that START in [s, e) are this thread's
			if (s > 0 && base[s - 1] != '\n') {
				const char *nl = (const char *) memchr(base + s, '\n', fsize - s);
				if (!nl) return;
				s = (size_t) (nl - base) + 1;
			}
			while (s < e) {
				if (s == 0 || base[s] == '>') found[t].push_back(s);
				const char *nl = (const char *) memchr(base + s, '\n', fsize - s);
				if (!nl) break;
				s = (size_t) (nl - base) + 1;
			}
		};
		std::vector<std::thread> th;
		for (unsigned t = 1; t < T; t++) th.emplace_back(scan, t);
		scan(0);
		for (auto &x : th) x.join();
	}
	std::vector<Record> recs;
	{
		size_t total = 0;
		for (auto &v : found) total += v.size();
		recs.reserve(total + 1);
		if (total == 0) { Record r; r.head = 0; recs.push_back(r); } // empty file: one empty record
		for (auto &v : found) for (size_t h : v) { Record r; r.head = h; recs.push_back(r); }
	}
	const size_t R = recs.size();
	parallel_for(R, T, 64, [&](size_t i) {
		Record &r = recs[i];
		const char *nl = r.head < fsize ? (const char *) memchr(base + r.head, '\n', fsize - r.head) : nullptr;
		r.head_end = nl ? (size_t) (nl - base) : fsize;
		r.seq = nl ? r.head_end + 1 : fsize;
		r.seq_end = i + 1 < R ? recs[i + 1].head : fsize;
	});

	// 2. measure and filter
	const bool want_gaps = !elastic && gap_limit != 0;
	const size_t grain = R / (T * 8) + 1;
	parallel_for(R, T, grain, [&](size_t i) { measure(base, recs[i], want_gaps); });
	const size_t expected = recs[0].len;
	MsaHolder *h = new MsaHolder();
	size_t m = 0;
	auto header_name = [&](const Record &r) { // identifier.substr(1)
		return r.head_end > r.head ? std::string(base + r.head + 1, base + r.head_end) : std::string();
	};
	for (size_t i = 0; i < R; i++) {
		Record &r = recs[i];
		if (output_paths) h->ids.push_back(header_name(r));
		if (r.len != expected) {
			h->messages += "WARNING: length of the sequence \xE2\x80\x9C" + header_name(r) + "\xE2\x80\x9D does not match that of the first sequence; skipping. (" +
			               std::to_string(expected) + " vs. " + std::to_string(r.len) + ")\n";
		} else if (want_gaps && r.gaprun >= gap_limit) {
			h->messages += "NOTICE: Sequence \xE2\x80\x9C" + header_name(r) + "\xE2\x80\x9D contained a gap run with " + std::to_string(r.gaprun) + " characters.\n";
		} else {
			r.keep = true; r.row = m++;
		}
	}

	// 3. copy the kept records into the row matrix
	const size_t n = expected;
	h->rows.reset(new uint8_t[m * n + 1]);
	uint8_t *rows = h->rows.get();
	parallel_for(R, T, grain, [&](size_t i) { if (recs[i].keep) copy_record(base, recs[i], rows + recs[i].row * n); });
	if (mapped) munmap(const_cast<char *>(base), fsize);

	for (auto &s : h->ids) h->id_ptrs.push_back(s.c_str());
	out->rows = rows; out->m = m; out->n = n;
	out->ids = h->id_ptrs.empty() ? nullptr : h->id_ptrs.data(); out->n_ids = h->id_ptrs.size();
	out->messages = h->messages.c_str();
	out->opaque = h;
	return FBG_OK;
}

extern "C" void fbg_msa_free(fbg_msa *msa) {
	if (!msa || !msa->opaque) return;
	delete static_cast<MsaHolder *>(msa->opaque);
	memset(msa, 0, sizeof *msa);
}
