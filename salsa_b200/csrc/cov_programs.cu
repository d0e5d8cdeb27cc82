// cov_programs.cu -- the two coverage tools of the reference as library calls:
//
//   break_contigs_start -a BED -l LEN -s N                  (main of break_contigs_start.cpp:271-463)
//   break_contigs       -a BED -b BREAKS -l LEN -i N -s N   (main of break_contigs.cpp:162-358)
//
// Same flags (short, --long=value and --long value forms of cmdline.h), same
// stdout.  salsa_b200/bin/break_contigs{,_start} are thin main()s over
// salsa_break_contigs_run; run_pipeline.py calls it directly so that all steps
// share one CUDA context and the BED the re-projection left in HBM.  Host code
// only: the kernels are coverage.cu.  Line order of the output follows the
// lengths file (the reference's order is std::unordered_map iteration order,
// break_contigs_start.cpp:393).
#include <fcntl.h>
#include <sys/mman.h>
#include <sys/stat.h>
#include <unistd.h>

#include <algorithm>
#include <cstdlib>
#include <fstream>
#include <iostream>
#include <map>
#include <numeric>
#include <sstream>
#include <string>
#include <unordered_map>
#include <vector>

#include "common.cuh"

using namespace std;

#define CK(expr)                                                                       \
    do {                                                                               \
        int rc__ = (expr);                                                             \
        if (rc__ != 0) { fprintf(stderr, "%s: %s\n", #expr, salsa_last_error()); return 1; } \
    } while (0)
#define CU(expr)                                                                       \
    do {                                                                               \
        cudaError_t e__ = (expr);                                                      \
        if (e__ != cudaSuccess) { fprintf(stderr, "%s: %s\n", #expr, cudaGetErrorString(e__)); return 1; } \
    } while (0)

namespace {

struct Mapping {
    const char *p = NULL;
    size_t n = 0;
    int fd = -1;
    // a missing file reads as empty, like the reference's ifstream loop
    int open(const char *path)
    {
        fd = ::open(path, O_RDONLY);
        if (fd < 0) return 0;
        struct stat sb;
        fstat(fd, &sb);
        n = (size_t)sb.st_size;
        if (n) {
            p = (const char *)mmap(NULL, n, PROT_READ, MAP_PRIVATE, fd, 0);
            if (p == MAP_FAILED) { perror("mmap"); p = NULL; return 1; }
        }
        return 0;
    }
    ~Mapping()
    {
        if (p) munmap((void *)p, n);
        if (fd >= 0) close(fd);
    }
};

int parse_args(int argc, const char *const *argv, const map<char, string> &shorts, const vector<string> &required,
               map<string, string> &o)
{
    for (int i = 0; i < argc; i++) {
        string a = argv[i], name, val;
        bool have = false;
        if (a.size() > 2 && a[0] == '-' && a[1] == '-') {
            size_t eq = a.find('=');
            name = a.substr(2, eq == string::npos ? string::npos : eq - 2);
            if (eq != string::npos) { val = a.substr(eq + 1); have = true; }
        } else if (a.size() == 2 && a[0] == '-' && shorts.count(a[1])) {
            name = shorts.at(a[1]);
        } else {
            fprintf(stderr, "undefined option: %s\n", a.c_str());
            return 1;
        }
        if (!have) {
            if (i + 1 >= argc) { fprintf(stderr, "option needs value: --%s\n", name.c_str()); return 1; }
            val = argv[++i];
        }
        o[name] = val;
    }
    for (const string &r : required)
        if (!o.count(r)) { fprintf(stderr, "need option: --%s\n", r.c_str()); return 1; }
    return 0;
}

static int run(bool joins, map<string, string> &opt, const char *d_resident, size_t resident_bytes, ostream &cout)
{
    // ---- lengths: `iss >> a >> n`, later line wins (break_contigs_start.cpp:286-294)
    unordered_map<string, long> contig_length;
    vector<string> order;            // first appearance in the lengths file
    {
        ifstream lf(opt["contiglen"]);
        string line;
        while (getline(lf, line)) {
            istringstream iss(line);
            string a;
            long n = 0;
            if (!(iss >> a)) continue;
            iss >> n;
            if (!contig_length.count(a)) order.push_back(a);
            contig_length[a] = n;
        }
    }
    // ---- breakpoints (break_contigs.cpp:67-108)
    unordered_map<string, vector<long>> kept;      // kept joins per scaffold, file order
    if (joins) {
        ifstream bf(opt["breakpoints"]);
        string line;
        while (getline(bf, line)) {
            istringstream iss(line);
            string contig;
            long pos;
            iss >> contig;
            vector<long> br;
            long clen = contig_length.count(contig) ? contig_length[contig] : 0;
            while (iss >> pos)
                if (pos - 2000000 >= 0 && pos + 2000000 <= clen) br.push_back(pos);
            kept[contig] = br;
        }
    }
    // ---- contig table in byte order (ids = ranks)
    vector<string> names(order);
    sort(names.begin(), names.end());
    const int n = (int)names.size();
    string blob;
    vector<int64_t> off(n + 1, 0), len(n, 0);
    vector<uint32_t> flags(n, SALSA_CTG_HAS_LEN);
    vector<uint64_t> covoff(n, ~0ull);
    uint64_t ndiff = 0;
    for (int i = 0; i < n; i++) {
        blob += names[i];
        off[i + 1] = (int64_t)blob.size();
        len[i] = contig_length[names[i]];
        if (len[i] >= 1000000) {                       // :325 / :218
            covoff[i] = ndiff;
            ndiff += ((uint64_t)len[i] + 10 + 4095) & ~4095ull;
        }
    }
    vector<uint64_t> joff(n + 1, 0);
    vector<int64_t> jpos;
    if (joins) {
        for (int i = 0; i < n; i++) {
            auto it = kept.find(names[i]);
            if (it != kept.end()) {
                vector<long> s(it->second);
                sort(s.begin(), s.end());
                for (long p : s) jpos.push_back(p);
            }
            joff[i + 1] = jpos.size();
        }
    }
    // ---- BED: the file, or the copy the caller already holds in device memory
    size_t fsz = 0;
    const char *bed = NULL;
    Mapping mapping;
    if (d_resident) {
        fsz = resident_bytes;
    } else {
        if (mapping.open(opt["alignment"].c_str())) return 1;
        bed = mapping.p;
        fsz = mapping.n;
    }
    // `prev_line == ""` skips leading empty lines plus the first non-empty line (:315-319)
    size_t begin = 0;
    if (d_resident) {
        vector<char> head(65536);
        for (bool more = true; more && begin < fsz;) {
            size_t k = min(head.size(), fsz - begin);
            CU(cudaMemcpy(head.data(), d_resident + begin, k, cudaMemcpyDeviceToHost));
            size_t j = 0;
            while (j < k && head[j] == '\n') j++;
            begin += j;
            more = (j == k);
        }
    } else {
        while (begin < fsz && bed[begin] == '\n') begin++;
    }

    // SALSA_B200_DEVICE if set, else the caller's current device (one process per GPU: a rank of a multi-GPU run calls this
    // with its own device current; a stand-alone program starts on device 0)
    int device = -1;
    if (getenv("SALSA_B200_DEVICE")) device = atoi(getenv("SALSA_B200_DEVICE"));
    else if (cudaGetDevice(&device) != cudaSuccess) device = 0;
    salsa_ctx *ctx = NULL;
    CK(salsa_ctx_create(device, &ctx));
    uint8_t *h_touched = (uint8_t *)calloc((size_t)n + 1, 1);
    int32_t *d_diff = NULL;
    if (n > 0 && fsz > begin) {
        CK(salsa_names_build(ctx, blob.data(), off.data(), n, len.data(), flags.data()));
        uint64_t *d_covoff, *d_joff, *d_stats;
        int64_t *d_len, *d_jpos;
        uint8_t *d_touched;
        CU(cudaMalloc(&d_covoff, (size_t)n * 8));
        CU(cudaMalloc(&d_len, (size_t)n * 8));
        CU(cudaMalloc(&d_joff, ((size_t)n + 1) * 8));
        CU(cudaMalloc(&d_jpos, (jpos.size() + 1) * 8));
        CU(cudaMalloc(&d_touched, (size_t)n + 1));
        CU(cudaMalloc(&d_stats, 32));
        CU(cudaMalloc(&d_diff, (ndiff + 1) * 4));
        CU(cudaMemcpy(d_covoff, covoff.data(), (size_t)n * 8, cudaMemcpyHostToDevice));
        CU(cudaMemcpy(d_len, len.data(), (size_t)n * 8, cudaMemcpyHostToDevice));
        CU(cudaMemcpy(d_joff, joff.data(), ((size_t)n + 1) * 8, cudaMemcpyHostToDevice));
        if (!jpos.empty()) CU(cudaMemcpy(d_jpos, jpos.data(), jpos.size() * 8, cudaMemcpyHostToDevice));
        const size_t CHUNK = d_resident ? fsz + 1 : (size_t)256 << 20;     // resident: one launch over the whole BED
        char *d_stage = NULL;
        if (!d_resident) CU(cudaMalloc(&d_stage, CHUNK + 64));
        size_t limit = fsz;          // bytes of the file that count (shrinks on a malformed line)
        for (int attempt = 0; attempt < 2; attempt++) {
            CU(cudaMemset(d_diff, 0, (ndiff + 1) * 4));
            CU(cudaMemset(d_touched, 0, (size_t)n + 1));
            CK(salsa_ingest_stats_reset(d_stats, NULL));
            size_t pos = begin;
            while (pos < limit) {
                size_t end = min(pos + CHUNK, limit), next = limit;
                if (end < limit) {
                    const char *nl = (const char *)memrchr(bed + pos, '\n', end - pos);
                    const char *pl = nl && nl > bed + pos ? (const char *)memrchr(bed + pos, '\n', (size_t)(nl - (bed + pos))) : NULL;
                    if (!nl || !pl) { fprintf(stderr, "BED line longer than the staging chunk\n"); return 1; }
                    end = (size_t)(nl - bed) + 1;
                    next = (size_t)(pl - bed) + 1;
                }
                const char *d_chunk = d_resident ? d_resident + pos : d_stage;
                if (!d_resident) CU(cudaMemcpy(d_stage, bed + pos, end - pos, cudaMemcpyHostToDevice));
                if (joins)
                    CK(salsa_cov_scan_joins(ctx, d_chunk, end - pos, pos, pos == begin, d_covoff, d_len, d_joff, d_jpos, d_diff, d_touched, d_stats, NULL));
                else
                    CK(salsa_cov_scan_start(ctx, d_chunk, end - pos, pos, pos == begin, d_covoff, d_len, d_diff, d_touched, d_stats, NULL));
                CU(cudaDeviceSynchronize());
                pos = next;
            }
            salsa_ingest_stats st;
            CK(salsa_ingest_stats_read(d_stats, &st, NULL));
            if (st.err_code == SALSA_DATA_MALFORMED && attempt == 0) {
                limit = (size_t)st.err_offset;     // the reference's loop breaks at this line
                continue;
            }
            if (st.err_code != 0) {
                fprintf(stderr, "BED data error %d at byte offset %llu\n", st.err_code, (unsigned long long)st.err_offset);
                return st.err_code == SALSA_DATA_RANGE ? 134 : 1;     // the reference aborts on vector::at()
            }
            break;
        }
        CU(cudaMemcpy(h_touched, d_touched, (size_t)n, cudaMemcpyDeviceToHost));
        if (d_stage) CU(cudaFree(d_stage));
        for (void *p : {(void *)d_covoff, (void *)d_len, (void *)d_joff, (void *)d_jpos, (void *)d_touched, (void *)d_stats}) cudaFree(p);
    }
    if (!joins) cerr << "bedfile loaded" << endl;
    // ---- segments
    vector<salsa_cov_seg> segs;
    vector<int> seg_contig;
    vector<long> seg_pos;
    uint64_t ncov = 0;
    for (const string &nm : order) {
        int i = (int)(lower_bound(names.begin(), names.end(), nm) - names.begin());
        if (!h_touched[i]) continue;
        if (!joins) {
            long sz = len[i] + 10;
            salsa_cov_seg s{covoff[i], covoff[i], ~0ull, sz, sz / 100, 9 * sz / 10};
            segs.push_back(s);
            seg_contig.push_back(i);
            seg_pos.push_back(0);
        } else {
            auto it = kept.find(nm);
            if (it == kept.end()) continue;
            for (long pos : it->second) {
                long start = pos - 2000000, sz = 4000000;
                // start == 0 reads tmpcov[-1] in the reference (undefined); read as 0 here
                salsa_cov_seg s{covoff[i] + (uint64_t)start, ncov, start > 0 ? covoff[i] + (uint64_t)start - 1 : ~0ull, sz, sz / 3, 2 * sz / 3};
                ncov += (uint64_t)sz;
                segs.push_back(s);
                seg_contig.push_back(i);
                seg_pos.push_back(pos);
            }
        }
    }
    const int nseg = (int)segs.size();
    vector<int64_t> sum(nseg + 1), ks(nseg * 11 + 1), ke(nseg * 11 + 1);
    // joins: windows are materialised in batches that fit a bounded scratch area
    const int batch = joins ? 256 : max(nseg, 1);
    int32_t *d_cov = d_diff;
    if (joins && nseg) CU(cudaMalloc(&d_cov, (size_t)min(nseg, batch) * 4000000 * 4));
    for (int b0 = 0; b0 < nseg; b0 += batch) {
        int nb = min(batch, nseg - b0);
        vector<salsa_cov_seg> part(segs.begin() + b0, segs.begin() + b0 + nb);
        if (joins) { uint64_t o = 0; for (auto &s : part) { s.dst_off = o; o += (uint64_t)s.len; } }
        size_t sb = salsa_cov_scratch_bytes(part.data(), nb);
        void *d_scratch;
        CU(cudaMalloc(&d_scratch, sb));
        CK(salsa_cov_analyse(d_diff, d_cov, part.data(), nb, d_scratch, sb, &sum[b0], &ks[(size_t)b0 * 11], &ke[(size_t)b0 * 11], NULL));
        CU(cudaFree(d_scratch));
    }
    // ---- decisions and output
    if (!joins) {
        for (int k = 0; k < nseg; k++) {
            long sz = segs[k].len;
            vector<long> positions;
            for (int q = 0; q < 11; q++) {
                long start = ks[(size_t)k * 11 + q], end = ke[(size_t)k * 11 + q];
                if (start >= sz / 100 + 5000 && end <= 9 * sz / 10 - 5000) positions.push_back((start + end) / 2);   // :450-453
            }
            if (!positions.empty()) {
                long consensus = std::accumulate(positions.begin(), positions.end(), 0.0) / positions.size();       // :456-460
                cout << names[seg_contig[k]] << "\t" << consensus << endl;
            }
        }
    } else {
        int total = 0, suspicious = 0;
        int k = 0;
        for (const string &nm : order) {
            int i = (int)(lower_bound(names.begin(), names.end(), nm) - names.begin());
            if (!h_touched[i]) continue;
            auto it = kept.find(nm);
            if (it == kept.end()) continue;                       // :264
            cout << nm;
            for (size_t j = 0; j < it->second.size(); j++, k++) {
                long pos = seg_pos[k], start_pos = pos - 2000000;
                int votes = 0;
                for (int q = 0; q < 11; q++) {
                    long start = ks[(size_t)k * 11 + q], end = ke[(size_t)k * 11 + q];
                    if (pos >= start + start_pos && pos <= end + start_pos) votes++;                                // :337-340
                }
                if (votes >= 8) { suspicious++; cout << "\t" << pos; }                                              // :344-348
                total++;
            }
            cout << endl;
        }
        cout << "Total Joins = " << total << endl;
        cout << "Suspicious Joins = " << suspicious << endl;
        cout << "Percent Suspicious Joins = " << suspicious * 100.0 / total << endl;
    }
    if (joins && d_cov != d_diff) cudaFree(d_cov);
    cudaFree(d_diff);
    free(h_touched);
    salsa_ctx_destroy(ctx);
    return 0;
}

}   // namespace

extern "C" int salsa_break_contigs_run(int32_t joins, int32_t argc, const char *const *argv, const void *d_bed,
                                       size_t nbytes, const char *out_path)
{
    map<string, string> opt;
    int rc;
    if (joins)
        rc = parse_args(argc, argv, {{'a', "alignment"}, {'b', "breakpoints"}, {'s', "min_size"}, {'l', "contiglen"}, {'i', "iteration"}},
                        {"alignment", "breakpoints", "min_size", "contiglen", "iteration"}, opt);
    else
        rc = parse_args(argc, argv, {{'a', "alignment"}, {'s', "min_size"}, {'l', "contiglen"}}, {"alignment", "min_size", "contiglen"}, opt);
    if (rc) return rc;
    ostringstream out;
    rc = run(joins != 0, opt, (const char *)d_bed, nbytes, out);
    const string text = out.str();
    if (out_path) {
        FILE *f = fopen(out_path, "wb");
        if (!f) { perror(out_path); return 1; }
        fwrite(text.data(), 1, text.size(), f);
        fclose(f);
    } else {
        fwrite(text.data(), 1, text.size(), stdout);
        fflush(stdout);
    }
    return rc;
}
