// kreeq-decompressor: reads the .bkwig files `kreeq validate -o x.bkwig` writes (writer: cli/writers.cpp,
// src/kreeq-output.cpp:305-399) -- `inflate` prints the whole file as .kwig text (or, with --expand, as the
// k-wide table of the .bed writer), `lookup` prints the triples of coordinate ranges.  Behaviour, options and
// stdout follow the reference tool (src/decompressor.cpp:76-618) including its coordinate arithmetic: ranges
// are 1-based inclusive on input, a range that reaches the end of its component is clipped and then read from
// the start of the data section (decompressor.cpp:137-150), and a header the index does not hold is fatal.
// One case is NOT reproducible: `lookup --expand` within k positions of the start of the data makes the
// reference seek to a wrapped-around offset and print an uninitialised array (decompressor.cpp:175-199); this
// tool prints zeros there.
// Host-only: nothing here touches the GPU or the k-mer table.
#include <getopt.h>
#include <unistd.h>

#include <algorithm>
#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <fstream>
#include <iostream>
#include <sstream>
#include <string>
#include <unordered_map>
#include <vector>

namespace {

struct Component { uint64_t byte_pos, abs_pos, len; uint8_t step; };   // BkwigIndexComponent, decompressor.cpp:34-39

struct Index {                                                          // BkwigIndex, decompressor.cpp:41-56
    std::unordered_map<std::string, std::vector<Component>> paths;
    std::vector<std::string> order;
    uint64_t bytes = 0;
    uint8_t k = 0;
};

struct Ranges {                                                         // BedCoordinates, gfalibs/include/bed.h:9-31
    std::unordered_map<std::string, std::vector<std::pair<uint64_t, uint64_t>>> by_header;
    std::vector<std::string> headers;
    void push(const std::string &h, uint64_t s, uint64_t e) {
        auto it = by_header.find(h);
        if (it == by_header.end()) { headers.push_back(h); by_header[h] = {{s, e}}; }
        else it->second.push_back({s, e});
    }
};

template <typename T> bool get(std::ifstream &f, T &v) { return (bool)f.read(reinterpret_cast<char *>(&v), sizeof(T)); }

// nPaths; per path: u16 name length, name, u32 nComponents; per component: u64 absPos, u64 len, u8 step.
// Data offsets run on across paths (decompressor.cpp:76-113).
void read_index(std::ifstream &f, Index &ix) {
    uint64_t byte_pos = 0;
    uint32_t npaths = 0;
    get(f, npaths); ix.bytes += 4;
    for (uint32_t i = 0; i < npaths && f; ++i) {
        uint16_t hs = 0; get(f, hs); ix.bytes += 2;
        std::string name(hs, '\0');
        f.read(&name[0], hs); ix.bytes += hs;
        uint32_t nc = 0; get(f, nc); ix.bytes += 4;
        std::vector<Component> comps;
        for (uint32_t e = 0; e < nc && f; ++e) {
            uint64_t abs_pos = 0, len = 0; uint8_t step = 0;
            get(f, abs_pos); get(f, len); get(f, step); ix.bytes += 17;
            comps.push_back({byte_pos, abs_pos, len, step});
            byte_pos += 12 * len;
        }
        ix.paths[name] = comps;
        ix.order.push_back(name);
    }
}

void print_triples(std::ifstream &f, uint64_t len) {                    // "cov,F,B\n" per base
    std::vector<uint32_t> v(len * 3, 0);
    f.read(reinterpret_cast<char *>(v.data()), (std::streamsize)(len * 12));
    std::string os;
    os.reserve(len * 12);
    for (uint64_t i = 0; i < len * 3; ++i) {
        os += std::to_string(v[i]);
        os += i % 3 < 2 ? ',' : '\n';
    }
    std::cout.write(os.data(), (std::streamsize)os.size());
}

// the .bed writer's row: name,pos, then the last k coverages, forward and backward edge supports
struct Window {
    uint8_t k; std::vector<uint32_t> cov, fw, bw;
    explicit Window(uint8_t kk) : k(kk), cov(kk ? kk - 1 : 0, 0), fw(cov), bw(cov) {}
    void prefill(const uint32_t *v) {
        cov.push_back(v[0]); fw.push_back(v[1]); bw.push_back(v[2]);
        cov.erase(cov.begin()); fw.erase(fw.begin()); bw.erase(bw.begin());
    }
    void row(const std::string &name, uint64_t pos, const uint32_t *v) {
        std::cout << name << ',' << pos << ',';
        std::vector<uint32_t> *cols[3] = {&cov, &fw, &bw};
        for (int c = 0; c < 3; ++c) {
            cols[c]->push_back(v[c]);
            for (uint8_t i = 0; i < k; ++i) {
                std::cout << std::to_string((*cols[c])[i]);
                if (i < k - 1) std::cout << ',';
            }
            if (c < 2) std::cout << ',';
        }
        std::cout << "\n";
        for (int c = 0; c < 3; ++c) cols[c]->erase(cols[c]->begin());
    }
};

void lookup(std::ifstream &f, const Index &ix, const std::string &name, const std::vector<std::pair<uint64_t, uint64_t>> &ranges,
            uint32_t span, bool expand) {                                // decompressor.cpp:115-249
    auto got = ix.paths.find(name);
    if (got == ix.paths.end()) {
        fprintf(stderr, "Could not find header (%s) Exiting.\n", name.c_str());
        exit(EXIT_FAILURE);
    }
    for (const auto &r : ranges) {
        uint64_t start = r.first - span - 1, end = r.second + span - 1;
        const uint64_t init = 1 + ix.bytes;
        uint64_t offset = init;
        for (const Component &c : got->second) {
            if (!(start >= c.abs_pos && start < c.abs_pos + c.len)) continue;
            if (end > c.abs_pos + c.len) end = c.abs_pos + c.len;
            else if (c.abs_pos + c.len > end) { offset += c.byte_pos + (start - c.abs_pos) * 12; break; }
        }
        uint64_t pos = start; const uint64_t len = end - start;
        if (!expand) {
            f.clear(); f.seekg((std::streamoff)offset);
            std::cout << name << ":" << start + 1 << "-" << end + 1 << std::endl;
            print_triples(f, len);
        } else {
            const uint8_t k = ix.k; uint8_t p = k;
            offset -= (uint64_t)k * 12;
            if (offset < init) { offset = init; p = (uint8_t)(k - pos); }
            f.clear(); f.seekg((std::streamoff)offset);
            Window w(k);
            uint32_t v[3] = {0, 0, 0};
            for (uint8_t i = 0; i < p; ++i) { f.read(reinterpret_cast<char *>(v), 12); w.prefill(v); }
            for (uint64_t i = 0; i < len; ++i) { f.read(reinterpret_cast<char *>(v), 12); w.row(name, pos++, v); }
        }
        std::cout << std::endl;
    }
}

bool is_number(const std::string &s) { return std::all_of(s.begin(), s.end(), [](char c) { return isdigit((unsigned char)c) != 0; }); }

// header[:begin-end], parsed from the back so that ':' and '-' may occur in the header (functions.h:697-724)
void parse_coordinate(std::string in, std::string &header, uint64_t &b, uint64_t &e) {
    std::reverse(in.begin(), in.end());
    const size_t dash = in.find('-'), colon = in.find(':');
    std::string sb = in.substr(dash + 1, colon - dash - 1), se = in.substr(0, dash);
    b = e = 0;
    if (is_number(se) && is_number(sb)) {
        header = in.substr(colon + 1);
        std::reverse(sb.begin(), sb.end()); std::reverse(se.begin(), se.end());
        b = std::stoull(sb); e = std::stoull(se);
    } else {
        header = in;
    }
    std::reverse(header.begin(), header.end());
}

void must_exist(const char *path) {                                    // ifFileExists, functions.h:260-272
    if (access(path, F_OK)) { std::cout << "Error - file does not exist: " << path << std::endl; exit(1); }
}

void usage() {
    printf("decompressor [mode]\n-h for additional help.\n");
    printf("\nModes:\n");
    printf("inflate\n");
    printf("lookup\n");
    exit(0);
}

}  // namespace

int main(int argc, char **argv) {
    if (argc < 2) usage();
    int expand = 0, verbose = 0, cmd = 0;
    int mode;
    if (!strcmp(argv[1], "inflate")) mode = 0;
    else if (!strcmp(argv[1], "lookup")) mode = 1;
    else { fprintf(stderr, "Unrecognized mode: %s\n", argv[1]); usage(); return 0; }
    const option longopts[] = {
        {"input-file", required_argument, 0, 'i'}, {"coordinate-file", required_argument, 0, 'c'}, {"span", required_argument, 0, 's'},
        {"out-format", required_argument, 0, 'o'}, {"max-memory", required_argument, 0, 'm'}, {"threads", required_argument, 0, 'j'},
        {"expand", no_argument, &expand, 1}, {"verbose", no_argument, &verbose, 1}, {"cmd", no_argument, &cmd, 1},
        {"version", no_argument, 0, 'v'}, {"help", no_argument, 0, 'h'}, {0, 0, 0, 0}};
    std::string input, coord_file;
    uint32_t span = 0;
    Ranges ranges;
    unsigned positional = 1;
    for (;;) {
        int oi = 0;
        const int c = getopt_long(argc, argv, mode == 0 ? "-:i:o:m:j:v:h" : "-:i:c:s:o:m:j:v:h", longopts, &oi);
        if (c == -1) break;
        switch (c) {
            case ':': fprintf(stderr, "option -%c is missing a required argument\n", optopt); return EXIT_FAILURE;
            case 0: break;
            case 'i': must_exist(optarg); input = optarg; break;
            case 'c': if (mode == 1) { must_exist(optarg); coord_file = optarg; } break;
            case 's': span = (uint32_t)atoi(optarg); break;
            case 'o': case 'm': case 'j': break;
            case 'v': printf("kreeq v%s\n", "0.0.1"); printf("Giulio Formenti giulio.formenti@gmail.com\n"); exit(0);
            case 'h':
                if (mode == 0) {
                    printf("kreeq inflate\n\nOptions:\n\t-i --input-file kreeq input.\n\t-o --out-format supported extensions:\n");
                } else {
                    printf("kreeq lookup [header[:start-end]]\n\nOptions:\n\t-i --input-file kreeq input.\n\t-c --coordinate-file sequence coordinates to extract.\n"
                           "\t-s --span <int> print context before and after coordinate positions.\n");
                }
                printf("\t-m --max-memory use at most this amount of memory (in Gb, default: 0.5 of max).\n\t-j --threads <n> numbers of threads (default: max).\n"
                       "\t-v --version software version.\n\t--cmd print $0 to stdout.\n");
                exit(0);
            default:                                    // positional: the mode word, then (lookup) one coordinate
                if (mode == 1) {
                    if (positional == 2) { std::string h; uint64_t b, e; parse_coordinate(optarg, h, b, e); ranges.push(h, b, e); }
                    else if (positional > 2) { printf("Error: too many positional arguments (%s).\n", optarg); exit(1); }
                    ++positional;
                }
                break;
        }
    }
    if (cmd) { for (int i = 0; i < argc; ++i) printf("%s ", argv[i]); printf("\n"); }
    std::ifstream f(input, std::ios::in | std::ios::binary);
    Index ix;
    get(f, ix.k);
    if (!expand) std::cout << std::to_string(ix.k) << "\n";
    read_index(f, ix);
    if (mode == 0) {                                    // decompressor.cpp:482-556
        for (const std::string &name : ix.order) {
            if (!f && f.peek() == EOF) { std::cout << "Error: file truncated" << std::endl; exit(EXIT_FAILURE); }
            for (const Component &c : ix.paths[name]) {
                if (!expand) {
                    std::cout << "fixedStep chrom=" << name << " start=" << std::to_string(c.abs_pos) << " step=" << std::to_string(c.step) << std::endl;
                    print_triples(f, c.len);
                } else {
                    Window w(ix.k);
                    uint64_t pos = c.abs_pos; uint32_t v[3] = {0, 0, 0};
                    for (uint64_t i = 0; i < c.len; ++i) { f.read(reinterpret_cast<char *>(v), 12); w.row(name, pos++, v); }
                }
            }
        }
        return EXIT_SUCCESS;
    }
    if (!coord_file.empty()) {                           // decompressor.cpp:560-572: "header begin end" per line
        std::ifstream bed(coord_file);
        std::string line, h; uint64_t b = 0, e = 0;
        while (getline(bed, line)) { std::istringstream iss(line); iss >> h >> b >> e; ranges.push(h, b, e); }
    }
    for (const std::string &h : ranges.headers) lookup(f, ix, h, ranges.by_header[h], span, expand != 0);
    return EXIT_SUCCESS;
}
