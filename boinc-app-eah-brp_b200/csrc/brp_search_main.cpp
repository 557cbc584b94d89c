// brp_b200_search: command-line driver over the BATCHED API of libbrp_b200.so.
//
// Same command line, input formats, candidate file and checkpoint file as the reference's standalone driver
// (demod_binary.c MAIN: options :217-470, work-unit :655-842, loop :1180-1443, results :1502-1686, checkpoint
// :1743-1783 / resume :560-652), but the template loop is brp_search_run() over whole chunks of the bank instead of
// one run_resampling/run_fft/run_harmonic_summing round trip per template.  (The unmodified reference driver linked
// against the eleven legacy symbols is the other way to run this library; see INTEGRATION.md.)
//
// Several GPUs of one box (--devices 0-7): one handle per device in this one process, every round of the loop hands each
// device a contiguous piece of the next stretch of the bank (so the searched templates are always a prefix of the bank and
// checkpoints keep their meaning), all devices run concurrently, and the per-device hit lists (a few KB each) come back
// by plain device-to-host copies and are merged on the host by brp_toplist_from_hits -- concatenation is exact, global
// template indices make "earliest template" well defined.  No NCCL here: inside one process there is nothing to
// exchange between the devices themselves.
//
// Several work-units in one invocation (-i a -o a.cand -i b -o b.cand ...): the passes of the reference's BOINC wrapper
// (erp_boinc_wrapper.cpp:430-474) -- each input is searched into the output of the same position, a pass whose output file
// exists already is skipped, and the checkpoint file is deleted after every completed pass so that the next one does not pick
// it up.  With --devices the passes are dealt out over the devices, one work-unit per device at a time (short work-units do not
// need more than one GPU each, and nothing is exchanged at all); each pass then keeps its own checkpoint file "<name>.<pass>".
//
// There is no CPU path in here: every stage of the search runs in the CUDA kernels of the library, and the program
// fails with the library's RADPUL_* code when no device is usable.
#include <stdio.h>
#include <stdlib.h>
#include <string.h>
#include <time.h>

#include <algorithm>
#include <string>
#include <thread>
#include <vector>

#include "../../include/brp_b200.h"

namespace {

struct Options {
    const char *input = nullptr, *output = nullptr, *bank = nullptr, *checkpoint = nullptr, *zaplist = nullptr;
    std::vector<const char *> inputs, outputs; /* every -i / -o given, in order: the passes */
    int white = 0, device = 0, debug = 0;
    float f0 = 250.0f, padding = 1.0f, fA = 0.04f; /* defaults of demod_binary.c:204-215 */
    unsigned window = 1000;
    unsigned chunk = 2048;            /* templates per brp_search_run between checkpoint opportunities */
    double checkpoint_period_s = 60;  /* wall-clock seconds between checkpoints (0 = after every chunk) */
    long max_templates = -1;          /* stop (with a checkpoint) after this many templates of this invocation */
    unsigned batch = 64;              /* templates per kernel launch */
    std::vector<int> devices;         /* --devices: several GPUs of this box (empty = the one of -D) */
};

/* "0-3", "0,2,5", "0-1,4": device ordinals in the order given (an ordinal may repeat: several handles on one GPU) */
bool parse_devices(const char *v, std::vector<int> &out)
{
    out.clear();
    const char *p = v;
    while (*p) {
        char *end = nullptr;
        long a = strtol(p, &end, 10);
        if (end == p || a < 0) return false;
        long b = a;
        if (*end == '-') {
            p = end + 1;
            b = strtol(p, &end, 10);
            if (end == p || b < a) return false;
        }
        for (long d = a; d <= b; d++) out.push_back((int)d);
        if (*end == ',') end++;
        else if (*end) return false;
        p = end;
    }
    return !out.empty() && out.size() <= 64;
}

void usage(const char *argv0)
{
    printf("Usage: %s [options]\n"
           " -i, --input_file       string  work-unit (.bin4 or .binary, zlib or plain); may be repeated, with as many -o\n"
           " -o, --output_file      string  candidate file (a pass whose output exists already is skipped when there are several)\n"
           " -t, --template_bank    string  template bank (P tau Psi0 per line)\n"
           " -c, --checkpoint_file  string  checkpoint file (read on start if present)\n"
           " -l, --zaplist_file     string  zap list: \"fmin fmax\" in Hz per line (required with -W)\n"
           " -f, --f0               float   maximum signal frequency in Hz (default 250)\n"
           " -A, --false_alarm      float   false alarm probability (default 0.04)\n"
           " -P, --padding          float   padding factor (default 1.0)\n"
           " -W, --whitening                whiten the series first\n"
           " -B, --box              int     running-median window (default 1000)\n"
           " -D, --device           int     CUDA device (default 0)\n"
           "     --devices          list    several GPUs of this box, e.g. 0-7 or 0,2,4: the bank is split over them\n"
           "                                (several work-units: one work-unit per device at a time instead)\n"
           " -z, --debug                    more output\n"
           "     --chunk            int     templates per enqueue between checkpoint opportunities (default 2048)\n"
           "     --batch            int     templates per kernel launch (default 64)\n"
           "     --checkpoint_period float  seconds between checkpoints (default 60; 0 = every chunk)\n"
           "     --max_templates    int     stop after this many templates, leaving a checkpoint\n",
           argv0);
}

bool is(const char *a, const char *s, const char *l) { return strcmp(a, s) == 0 || strcmp(a, l) == 0; }

int parse(int argc, char **argv, Options &o)
{
    for (int i = 1; i < argc;) {
        const char *a = argv[i];
        const bool has_val = i + 1 < argc;
        const char *v = has_val ? argv[i + 1] : "";
        if (is(a, "-W", "--whitening")) { o.white = 1; i++; continue; }
        if (is(a, "-z", "--debug")) { o.debug = 1; i++; continue; }
        if (is(a, "-h", "--help")) { usage(argv[0]); return RADPUL_EMISC; }
        if (!has_val) {
            fprintf(stderr, "Option %s needs a value.\n", a);
            return RADPUL_EMISC;
        }
        if (is(a, "-P", "--padding")) {
            if (atof(v) < 1.0 || atof(v) > 10.0) {
                fprintf(stderr, "Nonsense value: padding factor %g.\n", atof(v));
                return RADPUL_EVAL;
            }
            o.padding = atof(v);
        } else if (is(a, "-B", "--box")) {
            if (atoi(v) < 0 || atoi(v) > 250000) {
                fprintf(stderr, "Nonsense value: window size for running median %d.\n", atoi(v));
                return RADPUL_EVAL;
            }
            o.window = atoi(v);
        } else if (is(a, "-f", "--f0")) {
            if (atof(v) < 0.0 || atof(v) > 16.0e3) {
                fprintf(stderr, "Nonsense value: upper limit for search frequency %g.\n", atof(v));
                return RADPUL_EVAL;
            }
            o.f0 = atof(v);
        } else if (is(a, "-A", "--false_alarm")) {
            if (atof(v) < 0.0 || atof(v) > 1.0) {
                fprintf(stderr, "Nonsense value: false alarm rate %g.\n", atof(v));
                return RADPUL_EVAL;
            }
            o.fA = atof(v);
        } else if (is(a, "-i", "--input_file")) {
            if (!strstr(v, ".binary") && !strstr(v, ".bin4")) {
                fprintf(stderr, "Unknown file format (extension) for input file: %s\n", v);
                return RADPUL_EFILE;
            }
            o.input = v;
            o.inputs.push_back(v);
        } else if (is(a, "-o", "--output_file")) {
            o.output = v;
            o.outputs.push_back(v);
        }
        else if (is(a, "-c", "--checkpoint_file")) o.checkpoint = v;
        else if (is(a, "-t", "--template_bank")) o.bank = v;
        else if (is(a, "-l", "--zaplist_file")) o.zaplist = v;
        else if (is(a, "-D", "--device")) {
            if (atoi(v) < 0) {
                fprintf(stderr, "Nonsense value: device %d.\n", atoi(v));
                return RADPUL_EVAL;
            }
            o.device = atoi(v);
        } else if (strcmp(a, "--devices") == 0) {
            if (!parse_devices(v, o.devices)) {
                fprintf(stderr, "Nonsense value: device list %s.\n", v);
                return RADPUL_EVAL;
            }
        } else if (strcmp(a, "--chunk") == 0) o.chunk = std::max(1, atoi(v));
        else if (strcmp(a, "--batch") == 0) o.batch = std::max(1, atoi(v));
        else if (strcmp(a, "--checkpoint_period") == 0) o.checkpoint_period_s = atof(v);
        else if (strcmp(a, "--max_templates") == 0) o.max_templates = atol(v);
        else {
            fprintf(stderr, "Unknown option: %s\n", a);
            usage(argv[0]);
            return RADPUL_EMISC;
        }
        i += 2;
    }
    if (!o.input || !o.output || !o.bank) {
        fprintf(stderr, "Input file, output file and template bank are required.\n");
        return RADPUL_EMISC;
    }
    if (o.inputs.size() != o.outputs.size()) { /* erp_boinc_wrapper.cpp:412-416 */
        fprintf(stderr, "Number of input files (%zu) and output files (%zu) differ.\n", o.inputs.size(), o.outputs.size());
        return RADPUL_EMISC;
    }
    return 0;
}

double now_s()
{
    struct timespec ts;
    clock_gettime(CLOCK_MONOTONIC, &ts);
    return ts.tv_sec + 1e-9 * ts.tv_nsec;
}

}  // namespace

#define CHECK(call)                                                              \
    do {                                                                         \
        int rc_ = (call);                                                        \
        if (rc_ != 0) {                                                          \
            fprintf(stderr, "%s failed with code %d (%s:%d)\n", #call, rc_, __FILE__, __LINE__); \
            return rc_;                                                          \
        }                                                                        \
    } while (0)

/* one work-unit: o.input -> o.output on o.devices (the whole of demod_binary.c MAIN behind the option parsing) */
static int run_pass(Options o)
{
    /* ---- resume: demod_binary.c:560-652 ---- */
    uint32_t done = 0;
    bool resumed = false;
    std::vector<brp_cand> carried(BRP_N_CAND);
    memset(carried.data(), 0, sizeof(brp_cand) * BRP_N_CAND);
    if (o.checkpoint) {
        char original[256] = {0};
        FILE *probe = fopen(o.checkpoint, "rb");
        if (probe) {
            fclose(probe);
            CHECK(brp_checkpoint_read(o.checkpoint, &done, original, carried.data()));
            if (strcmp(original, o.input) != 0) {
                fprintf(stderr, "Input file on command line %s doesn't agree with input file %s from checkpoint header.\n",
                        o.input, original);
                return RADPUL_EVAL;
            }
            resumed = true;
            fprintf(stderr, "Continuing work on %s at template no. %u\n", original, done);
        }
    }

    /* ---- work-unit and bank ---- */
    float *series = nullptr;
    uint32_t nu = 0;
    double tsample_us = 0;
    CHECK(brp_wu_read(o.input, &series, &nu, &tsample_us));
    double *bank = nullptr;
    uint32_t n_templates = 0;
    CHECK(brp_bank_read(o.bank, &bank, &n_templates));
    if (done > n_templates) {
        fprintf(stderr, "Checkpoint is ahead of the template bank (%u > %u).\n", done, n_templates);
        return RADPUL_EVAL;
    }
    brp_geometry g;
    CHECK(brp_derive_geometry(nu, tsample_us, o.padding, o.f0, o.window, &g));

    /* ---- whitening, once per work-unit (demod_binary.c:858-1079) ---- */
    if (o.white) {
        /* the reference opens the zap list unconditionally once -W is given (:982-989) */
        if (!o.zaplist) {
            fprintf(stderr, "Couldn't open zaplist file: (none given; -W needs -l)\n");
            return RADPUL_EFILE;
        }
        double *zap = nullptr;
        uint32_t n_zap = 0;
        CHECK(brp_zaplist_read(o.zaplist, &zap, &n_zap));
        float ms = 0;
        CHECK(brp_whiten_series_zap(o.device, series, nu, o.padding, o.window, tsample_us, zap, n_zap, &ms));
        brp_zaplist_free(zap);
        if (o.debug) fprintf(stderr, "Whitening: %.2f ms on the device, %u zap ranges\n", ms, n_zap);
    }

    /* ---- the template loop, chunk by chunk; one handle per device ---- */
    if (o.devices.empty()) o.devices.push_back(o.device);
    const size_t ndev = o.devices.size();
    std::vector<brp_search *> handles(ndev, nullptr);
    for (size_t d = 0; d < ndev; d++) {
        CHECK(brp_search_create(&handles[d], o.devices[d], &g, o.fA, o.batch));
        CHECK(brp_search_set_series(handles[d], series));
    }
    for (size_t d = 0; d < ndev; d++) CHECK(brp_search_sync(handles[d])); /* the copies read `series` */
    brp_series_free(series);

    const size_t cap = 5u * (size_t)g.fundamental_idx_hi + 1;
    std::vector<brp_hit> hits(cap * ndev);
    std::vector<brp_cand> fresh(BRP_N_CAND), merged(BRP_N_CAND);
    /* candidates of everything searched so far: checkpoint content merged with this invocation's device tables
     * (hit lists of the devices concatenated: exact, see brp_toplist_from_hits) */
    auto snapshot = [&](brp_cand *out) -> int {
        uint32_t n_hits = 0;
        for (size_t d = 0; d < ndev; d++) {
            uint32_t n_d = 0;
            int r = brp_search_collect(handles[d], hits.data() + n_hits, (uint32_t)cap, &n_d);
            if (r) return r;
            n_hits += n_d;
        }
        int r = brp_toplist_from_hits(hits.data(), n_hits, bank, n_templates, fresh.data(), nullptr);
        if (r) return r;
        if (!resumed) {
            memcpy(out, fresh.data(), sizeof(brp_cand) * BRP_N_CAND);
            return 0;
        }
        return brp_toplist_merge(carried.data(), fresh.data(), out);
    };

    const double t_start = now_s();
    double t_last_cp = t_start;
    uint32_t pos = done;
    long budget = o.max_templates;
    bool stopped_early = false, wrote_checkpoint = false;
    while (pos < n_templates) {
        /* one round: the next `chunk` templates per device, cut into contiguous pieces, all devices at once */
        uint32_t n = (uint32_t)std::min<uint64_t>((uint64_t)o.chunk * ndev, n_templates - pos);
        if (budget >= 0) {
            if (budget == 0) {
                stopped_early = true;
                break;
            }
            n = (uint32_t)std::min<long>(n, budget);
            budget -= n;
        }
        /* one host thread per device enqueues (and awaits) that device's piece: the first launches on a device load its kernels,
         * which takes milliseconds and would otherwise be paid device after device */
        std::vector<int> rcs(ndev, 0);
        auto piece = [&](size_t d) {
            const uint32_t lo = (uint32_t)((uint64_t)n * d / ndev), hi = (uint32_t)((uint64_t)n * (d + 1) / ndev);
            if (hi > lo) rcs[d] = brp_search_run(handles[d], bank + 3 * (size_t)(pos + lo), hi - lo, pos + lo); /* asynchronous */
            if (!rcs[d]) rcs[d] = brp_search_sync(handles[d]);
        };
        if (ndev == 1) {
            piece(0);
        } else {
            std::vector<std::thread> workers;
            for (size_t d = 0; d < ndev; d++) workers.emplace_back(piece, d);
            for (std::thread &w : workers) w.join();
        }
        for (size_t d = 0; d < ndev; d++) CHECK(rcs[d]);
        pos += n;
        const bool last = pos == n_templates;
        const bool due = o.checkpoint && !last && (now_s() - t_last_cp >= o.checkpoint_period_s || budget == 0);
        if (due) {
            CHECK(snapshot(merged.data()));
            CHECK(brp_checkpoint_write(o.checkpoint, pos, o.input, merged.data()));
            t_last_cp = now_s();
            wrote_checkpoint = true;
            if (o.debug) fprintf(stderr, "Checkpoint at template %u\n", pos);
        }
    }
    if (budget == 0 && pos < n_templates) stopped_early = true;
    const double t_loop = now_s() - t_start;
    if (o.debug)
        fprintf(stderr, "%u templates on %zu device(s) in %.3f s (%.0f templates/s)\n", pos - done, ndev, t_loop,
                (pos - done) / std::max(t_loop, 1e-9));

    if (stopped_early) {
        fprintf(stderr, "Stopped after %u of %u templates%s.\n", pos, n_templates, wrote_checkpoint ? "; checkpoint written" : "");
        for (brp_search *h : handles) brp_search_destroy(h);
        brp_bank_free(bank);
        return 0;
    }

    /* ---- results: demod_binary.c:1502-1686 ---- */
    CHECK(snapshot(merged.data()));
    if (o.checkpoint) CHECK(brp_checkpoint_write(o.checkpoint, n_templates, o.input, merged.data())); /* :1498 final checkpoint */
    CHECK(brp_finalize_candidates(merged.data(), nullptr));
    CHECK(brp_write_candidates(o.output, merged.data(), &g, tsample_us));
    for (brp_search *h : handles) CHECK(brp_search_destroy(h));
    brp_bank_free(bank);
    return 0;
}

int main(int argc, char **argv)
{
    Options o;
    int rc = parse(argc, argv, o);
    if (rc) return rc;
    const size_t passes = o.inputs.size();
    if (passes == 1) return run_pass(o);

    /* ---- several work-units: the wrapper's passes (erp_boinc_wrapper.cpp:430-474) ---- */
    if (o.devices.empty()) o.devices.push_back(o.device);
    const size_t ndev = std::min(o.devices.size(), passes);
    std::vector<int> rcs(passes, 0);
    std::vector<std::string> cp_names(passes);
    auto one = [&](size_t pass, int device) {
        Options po = o;
        po.input = o.inputs[pass];
        po.output = o.outputs[pass];
        po.device = device;
        po.devices.assign(1, device);
        if (o.checkpoint) {
            /* sequential passes share the one checkpoint file like the wrapper's; concurrent ones need one each */
            cp_names[pass] = ndev > 1 ? std::string(o.checkpoint) + "." + std::to_string(pass) : std::string(o.checkpoint);
            po.checkpoint = cp_names[pass].c_str();
        }
        if (FILE *fp = fopen(po.output, "r")) { /* :452-456 */
            fclose(fp);
            fprintf(stderr, "Output file: '%s' already exists - skipping pass\n", po.output);
            return;
        }
        rcs[pass] = run_pass(po);
        if (rcs[pass]) {
            fprintf(stderr, "Demodulation failed (error: %i)!\n", rcs[pass]);
            return;
        }
        /* a pass that stopped early (--max_templates) has no output yet and keeps its checkpoint */
        FILE *done = fopen(po.output, "r");
        if (done) fclose(done);
        if (done && po.checkpoint) remove(po.checkpoint); /* :466-467 */
    };
    if (ndev == 1) {
        for (size_t pass = 0; pass < passes; pass++) {
            one(pass, o.devices[0]);
            if (rcs[pass]) return rcs[pass]; /* the wrapper stops at the first failing pass */
        }
        return 0;
    }
    std::vector<std::thread> workers;
    for (size_t d = 0; d < ndev; d++)
        workers.emplace_back([&, d]() {
            for (size_t pass = d; pass < passes; pass += ndev) {
                one(pass, o.devices[d]);
                if (rcs[pass]) return;
            }
        });
    for (std::thread &w : workers) w.join();
    for (size_t pass = 0; pass < passes; pass++)
        if (rcs[pass]) return rcs[pass];
    return 0;
}
