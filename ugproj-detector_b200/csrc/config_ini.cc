// Reader for the [detection] (and [scan] frame_size) keys of the reference's conf/*.cfg files.
// Same keys, value syntax and defaults as Configuration::load (reference src/structure.cc:62-85,
// validators :258-300, path resolution :199-204), without Boost.program_options.
#include <cstdio>
#include <cstring>
#include <string>

#include "../../include/ugdet.h"

namespace {

std::string trim(const std::string& s) {
    size_t a = s.find_first_not_of(" \t\r\n"), b = s.find_last_not_of(" \t\r\n");
    return a == std::string::npos ? std::string() : s.substr(a, b - a + 1);
}

bool parse_triplet(const std::string& v, uint8_t out[3]) {      // "a,b,c"  (structure.cc:258-277)
    unsigned a, b, c;
    if (sscanf(v.c_str(), "%3u , %3u , %3u", &a, &b, &c) != 3 || a > 255 || b > 255 || c > 255) return false;
    out[0] = (uint8_t)a; out[1] = (uint8_t)b; out[2] = (uint8_t)c;
    return true;
}

}  // namespace

extern "C" void ugd_params_default(ugd_params* p) {
    if (!p) return;
    memset(p, 0, sizeof *p);
    p->scale_factor = 1.1;                 // structure.cc:83-85
    p->min_neighbors = 2;                  // detector.cc:29
    p->min_w = p->min_h = 30;              // detector.cc:31
    p->max_w = p->max_h = 0;
    p->skin_lower[0] = 0;   p->skin_lower[1] = 133; p->skin_lower[2] = 77;    // structure.cc:75-78
    p->skin_upper[0] = 255; p->skin_upper[1] = 173; p->skin_upper[2] = 127;   // structure.cc:79-82
    p->dense_stages = 0;
    p->max_raw_per_frame = 0;
}

extern "C" int ugd_config_load(const char* cfg_path, ugd_params* p, char* cascade_path, size_t cap,
                               int32_t* scan_w, int32_t* scan_h) {
    if (!cfg_path || !p) return UGD_ERR_ARG;
    FILE* f = fopen(cfg_path, "r");
    if (!f) return UGD_ERR_IO;
    ugd_params_default(p);
    if (scan_w) *scan_w = -1;              // structure.cc:64-65 default Size(-1,-1)
    if (scan_h) *scan_h = -1;
    std::string section, casc;
    char line[4096];
    int rc = UGD_OK;
    while (fgets(line, sizeof line, f)) {
        std::string s(line);
        size_t hash = s.find('#');
        if (hash != std::string::npos) s.erase(hash);
        s = trim(s);
        if (s.empty()) continue;
        if (s.front() == '[') {
            if (s.back() != ']') { rc = UGD_ERR_PARSE; break; }
            section = trim(s.substr(1, s.size() - 2));
            continue;
        }
        size_t eq = s.find('=');
        if (eq == std::string::npos) { rc = UGD_ERR_PARSE; break; }
        std::string key = section.empty() ? trim(s.substr(0, eq)) : section + "." + trim(s.substr(0, eq));
        std::string val = trim(s.substr(eq + 1));
        if (key == "detection.skin_lower") { if (!parse_triplet(val, p->skin_lower)) { rc = UGD_ERR_PARSE; break; } }
        else if (key == "detection.skin_upper") { if (!parse_triplet(val, p->skin_upper)) { rc = UGD_ERR_PARSE; break; } }
        else if (key == "detection.scale") {
            char* e = nullptr; double d = strtod(val.c_str(), &e);
            if (e == val.c_str() || !(d > 1.0)) { rc = UGD_ERR_PARSE; break; }
            p->scale_factor = d;
        } else if (key == "detection.cascade_classifier_filepath") casc = val;
        else if (key == "scan.frame_size") {                       // "WxH" (structure.cc:281-300)
            unsigned a, b;
            if (sscanf(val.c_str(), "%u x %u", &a, &b) != 2) { rc = UGD_ERR_PARSE; break; }
            if (scan_w) *scan_w = (int32_t)a;
            if (scan_h) *scan_h = (int32_t)b;
        }
        // every other section/key belongs to the tracker / clusterer and is ignored here
    }
    fclose(f);
    if (rc != UGD_OK) return rc;
    if (casc.empty()) return UGD_ERR_PARSE;        // the key is ->required() (structure.cc:73)
    if (casc[0] != '/') {                          // relative to the cfg's directory (structure.cc:199-204)
        std::string dir(cfg_path);
        size_t sl = dir.find_last_of('/');
        dir = sl == std::string::npos ? std::string(".") : dir.substr(0, sl);
        casc = dir + "/" + casc;
    }
    if (cascade_path) {
        if (casc.size() + 1 > cap) return UGD_ERR_ARG;
        memcpy(cascade_path, casc.c_str(), casc.size() + 1);
    }
    return UGD_OK;
}
