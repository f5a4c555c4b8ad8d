// hog_storage.cu -- host only: the "opencv-object-detector-hog" FileStorage node (XML or YAML by file extension),
// i.e. what cv::HOGDescriptor::write / read (modules/objdetect/src/hog.cpp:139-196) and cv::cuda::HOG's
// write / read (modules/cudaobjdetect/src/hog.cpp:599-679) exchange, including the SVMDetector sequence that the
// fork's GPU writer leaves out (its `--svm <file>` path, samples/gpu/tbd.cpp:417-427, needs it).
//
// Not a general FileStorage: it writes exactly the layout the reference writes (so the reference loads our files)
// and reads maps of scalars / numeric sequences in the XML and YAML forms the reference emits.
#include "../../include/hogb200.h"

#include <cctype>
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <map>
#include <string>
#include <vector>

namespace hogk { int storage_fail(int code, const std::string& msg); }

namespace {

typedef std::map<std::string, std::vector<std::string> > Node;     // key -> tokens (numbers as text)

bool ends_with(const std::string& s, const char* suf)
{
    const size_t n = strlen(suf);
    return s.size() >= n && strcasecmp(s.c_str() + s.size() - n, suf) == 0;
}

// FileStorage::getDefaultObjectName, modules/core/src/persistence_cpp.cpp:192-229
std::string default_object_name(const std::string& filename)
{
    size_t end = filename.size(), p = filename.size();
    while (p > 0 && filename[p - 1] != '\\' && filename[p - 1] != '/' && filename[p - 1] != ':') {
        p--;
        if (filename[p] == '.' && (end == filename.size() || filename.compare(end, 3, ".gz") == 0)) end = p;
    }
    if (p == end) return "";
    std::string name;
    if (!isalpha((unsigned char)filename[p]) && filename[p] != '_') name += '_';
    for (; p < end; p++) {
        const char c = filename[p];
        name += (isalnum((unsigned char)c) || c == '-' || c == '_') ? c : '_';
    }
    return name == "_" ? "unnamed" : name;
}

// icvDoubleToString / icvFloatToString, modules/core/src/persistence.cpp: integers print as "%d.", the rest as
// %.16e / %.8e (both round-trip)
std::string real_text(double v, bool is_float)
{
    char buf[64];
    if (std::isnan(v)) return ".Nan";
    if (std::isinf(v)) return v < 0 ? "-.Inf" : ".Inf";
    // `int ivalue = cvRound(value); if (ivalue == value)`: only values an int can hold print as integers
    if (v >= -2147483648.0 && v < 2147483648.0 && (double)std::lrint(v) == v) snprintf(buf, sizeof buf, "%ld.", std::lrint(v));
    else snprintf(buf, sizeof buf, is_float ? "%.8e" : "%.16e", v);
    return buf;
}

double text_real(const std::string& t)
{
    if (t == ".Nan" || t == ".NaN" || t == ".nan") return NAN;
    if (t == ".Inf" || t == ".inf" || t == "+.Inf") return INFINITY;
    if (t == "-.Inf" || t == "-.inf") return -INFINITY;
    return strtod(t.c_str(), nullptr);
}

struct Field { const char* key; std::vector<std::string> tok; bool seq; };

std::vector<Field> model_fields(const hogb200_model* m, const float* det, size_t n)
{
    auto I = [](int v) { return std::to_string(v); };
    std::vector<Field> f;
    const hogb200_geom& g = m->geom;
    f.push_back({"winSize", {I(g.win_w), I(g.win_h)}, true});
    f.push_back({"blockSize", {I(g.block_w), I(g.block_h)}, true});
    f.push_back({"blockStride", {I(g.block_stride_w), I(g.block_stride_h)}, true});
    f.push_back({"cellSize", {I(g.cell_w), I(g.cell_h)}, true});
    f.push_back({"nbins", {I(g.nbins)}, false});
    f.push_back({"derivAperture", {"1"}, false});
    // getWinSigma(): a negative sigma stands for (block_w + block_h) / 8 (hog.cpp:104-107)
    const double ws = m->win_sigma >= 0 ? m->win_sigma : (g.block_w + g.block_h) / 8.;
    f.push_back({"winSigma", {real_text(ws, false)}, false});
    f.push_back({"histogramNormType", {"0"}, false});
    f.push_back({"L2HysThreshold", {real_text(m->l2hys_threshold, false)}, false});
    f.push_back({"gammaCorrection", {I(m->gamma_correction ? 1 : 0)}, false});
    f.push_back({"nlevels", {I(m->nlevels)}, false});
    f.push_back({"signedGradient", {"0"}, false});
    if (det && n) {
        Field d{"SVMDetector", {}, true};
        d.tok.reserve(n);
        for (size_t i = 0; i < n; i++) d.tok.push_back(real_text(det[i], true));
        f.push_back(d);
    }
    return f;
}

// Line filling as the reference's emitters do it (icvYMLWrite / icvXMLWriteScalar, modules/core/src/persistence.cpp,
// wrap_margin 71): a value moves to the next line when it would end past column 71 and the line is not nearly empty.
bool wraps(size_t line_len, size_t tok_len, size_t indent)
{
    const size_t end = line_len + tok_len;
    return end > 71 && end - indent > 10;
}

void write_yaml(FILE* fp, const std::string& name, const std::vector<Field>& f)
{
    fprintf(fp, "%%YAML:1.0\n---\n%s: !!opencv-object-detector-hog\n", name.c_str());
    const size_t indent = 7;             // struct_indent of a flow sequence inside the object's map
    for (const Field& x : f) {
        if (!x.seq) { fprintf(fp, "   %s: %s\n", x.key, x.tok[0].c_str()); continue; }
        std::string line = std::string("   ") + x.key + ": [";
        for (size_t i = 0; i < x.tok.size(); i++) {
            if (i) line += ',';
            if (wraps(line.size(), x.tok[i].size(), indent)) { fprintf(fp, "%s\n", line.c_str()); line.assign(indent, ' '); }
            else line += ' ';
            line += x.tok[i];
        }
        fprintf(fp, "%s ]\n", line.c_str());
    }
}

void write_xml(FILE* fp, const std::string& name, const std::vector<Field>& f)
{
    fprintf(fp, "<?xml version=\"1.0\"?>\n<opencv_storage>\n<%s type_id=\"opencv-object-detector-hog\">", name.c_str());
    const size_t indent = 4;
    for (const Field& x : f) {
        if (!x.seq) { fprintf(fp, "\n  <%s>%s</%s>", x.key, x.tok[0].c_str(), x.key); continue; }
        fprintf(fp, "\n  <%s>\n", x.key);
        std::string line(indent, ' ');
        for (size_t i = 0; i < x.tok.size(); i++) {
            if (wraps(line.size(), x.tok[i].size(), indent)) { fprintf(fp, "%s\n", line.c_str()); line.assign(indent, ' '); }
            else if (line.size() > indent) line += ' ';
            line += x.tok[i];
        }
        fprintf(fp, "%s</%s>", line.c_str(), x.key);
    }
    fprintf(fp, "</%s>\n</opencv_storage>\n", name.c_str());     // the closing tag follows the last child directly
}

void split_numbers(const std::string& text, std::vector<std::string>& out)
{
    size_t i = 0;
    while (i < text.size()) {
        while (i < text.size() && (isspace((unsigned char)text[i]) || text[i] == ',' || text[i] == '[' || text[i] == ']')) i++;
        size_t j = i;
        while (j < text.size() && !isspace((unsigned char)text[j]) && text[j] != ',' && text[j] != ']') j++;
        if (j > i) out.push_back(text.substr(i, j - i));
        i = j;
    }
}

// ---- XML: <opencv_storage> <name ...> <key>numbers</key> ... </name> </opencv_storage>
bool read_xml(const std::string& s, const std::string& want, Node& node, std::string& err)
{
    size_t p = s.find("<opencv_storage>");
    if (p == std::string::npos) { err = "no <opencv_storage> element"; return false; }
    p += 16;
    int depth = 0;
    std::string cur_key;
    bool inside = false;
    while (p < s.size()) {
        const size_t lt = s.find('<', p);
        if (lt == std::string::npos) break;
        if (inside && depth == 2 && !cur_key.empty()) split_numbers(s.substr(p, lt - p), node[cur_key]);
        if (s.compare(lt, 4, "<!--") == 0) { const size_t e = s.find("-->", lt); if (e == std::string::npos) break; p = e + 3; continue; }
        const size_t gt = s.find('>', lt);
        if (gt == std::string::npos) { err = "unterminated tag"; return false; }
        std::string tag = s.substr(lt + 1, gt - lt - 1);
        p = gt + 1;
        if (tag.empty() || tag[0] == '?') continue;
        if (tag[0] == '/') {
            const std::string name = tag.substr(1);
            if (name == "opencv_storage") break;
            if (depth == 2) cur_key.clear();
            if (depth == 1 && inside) return true;
            depth--;
            continue;
        }
        const bool self_closing = tag[tag.size() - 1] == '/';
        const std::string name = tag.substr(0, tag.find_first_of(" \t\r\n/"));
        if (self_closing) continue;
        depth++;
        if (depth == 1) inside = want.empty() || name == want;
        else if (depth == 2 && inside) { cur_key = name; node[cur_key]; }
    }
    if (!inside) err = "object '" + want + "' not found";
    return inside && !node.empty();
}

// ---- YAML: "name: [!!type]" at column 0, then indented "key: scalar" / "key: [ flow sequence, possibly wrapped ]"
bool read_yaml(const std::string& s, const std::string& want, Node& node, std::string& err)
{
    std::vector<std::string> lines;
    for (size_t p = 0; p < s.size();) {
        size_t e = s.find('\n', p);
        if (e == std::string::npos) e = s.size();
        std::string l = s.substr(p, e - p);
        if (!l.empty() && l[l.size() - 1] == '\r') l.erase(l.size() - 1);
        lines.push_back(l);
        p = e + 1;
    }
    bool inside = false, found = false;
    std::string open_key;                 // a flow sequence still waiting for its ']'
    for (const std::string& l : lines) {
        if (l.empty() || l[0] == '%' || l[0] == '#' || l.compare(0, 3, "---") == 0 || l.compare(0, 3, "...") == 0) continue;
        if (!isspace((unsigned char)l[0])) {                        // top-level key
            if (inside) break;
            const size_t c = l.find(':');
            if (c == std::string::npos) continue;
            const std::string name = l.substr(0, c);
            inside = found = want.empty() || name == want;
            continue;
        }
        if (!inside) continue;
        if (!open_key.empty()) {
            split_numbers(l, node[open_key]);
            if (l.find(']') != std::string::npos) open_key.clear();
            continue;
        }
        size_t b = 0;
        while (b < l.size() && isspace((unsigned char)l[b])) b++;
        const size_t c = l.find(':', b);
        if (c == std::string::npos) continue;
        const std::string key = l.substr(b, c - b), val = l.substr(c + 1);
        split_numbers(val, node[key]);
        if (val.find('[') != std::string::npos && val.find(']') == std::string::npos) open_key = key;
    }
    if (!found) err = "object '" + want + "' not found";
    return found && !node.empty();
}

bool get_ints(const Node& n, const char* key, int* out, size_t cnt)
{
    Node::const_iterator it = n.find(key);
    if (it == n.end() || it->second.size() < cnt) return false;
    for (size_t i = 0; i < cnt; i++) out[i] = (int)std::lrint(text_real(it->second[i]));
    return true;
}

}  // namespace

extern "C" {

int hogb200_model_write(const char* filename, const char* objname, const hogb200_model* m, const float* detector, size_t n)
{
    if (!filename || !m) return hogk::storage_fail(HOGB200_BAD_ARG, "model_write: null argument");
    const std::string fn(filename);
    const bool xml = ends_with(fn, ".xml"), yml = ends_with(fn, ".yml") || ends_with(fn, ".yaml");
    if (!xml && !yml) return hogk::storage_fail(HOGB200_BAD_ARG, "model_write: the file name must end in .xml, .yml or .yaml");
    const std::string name = objname && *objname ? std::string(objname) : default_object_name(fn);
    if (name.empty()) return hogk::storage_fail(HOGB200_BAD_ARG, "model_write: invalid file name");
    FILE* fp = fopen(filename, "w");
    if (!fp) return hogk::storage_fail(HOGB200_BAD_ARG, "model_write: cannot open " + fn);
    const std::vector<Field> f = model_fields(m, detector, n);
    if (xml) write_xml(fp, name, f); else write_yaml(fp, name, f);
    const bool ok = fclose(fp) == 0;
    return ok ? HOGB200_OK : hogk::storage_fail(HOGB200_BAD_ARG, "model_write: write failed");
}

int hogb200_model_read(const char* filename, const char* objname, hogb200_model* m, float* detector, size_t capacity,
                       size_t* count)
{
    if (count) *count = 0;
    if (!filename || !m) return hogk::storage_fail(HOGB200_BAD_ARG, "model_read: null argument");
    FILE* fp = fopen(filename, "r");
    if (!fp) return hogk::storage_fail(HOGB200_BAD_ARG, std::string("model_read: cannot open ") + filename);
    std::string s;
    char buf[1 << 16];
    size_t got;
    while ((got = fread(buf, 1, sizeof buf, fp)) > 0) s.append(buf, got);
    fclose(fp);
    size_t b = 0;
    while (b < s.size() && isspace((unsigned char)s[b])) b++;
    Node node;
    std::string err;
    const std::string want = objname ? objname : "";
    const bool ok = s.compare(b, 5, "<?xml") == 0 || s.compare(b, 1, "<") == 0 ? read_xml(s, want, node, err)
                                                                             : read_yaml(s, want, node, err);
    if (!ok) return hogk::storage_fail(HOGB200_BAD_ARG, "model_read: " + (err.empty() ? std::string("no HOG node") : err));
    int v[2];
    hogb200_geom& g = m->geom;
    // hog.cpp:141 CV_Assert(!obj["winSize"].empty())
    if (!get_ints(node, "winSize", v, 2)) return hogk::storage_fail(HOGB200_BAD_ARG, "model_read: winSize missing");
    g.win_w = v[0]; g.win_h = v[1];
    if (!get_ints(node, "blockSize", v, 2)) return hogk::storage_fail(HOGB200_BAD_ARG, "model_read: blockSize missing");
    g.block_w = v[0]; g.block_h = v[1];
    if (!get_ints(node, "blockStride", v, 2)) return hogk::storage_fail(HOGB200_BAD_ARG, "model_read: blockStride missing");
    g.block_stride_w = v[0]; g.block_stride_h = v[1];
    if (!get_ints(node, "cellSize", v, 2)) return hogk::storage_fail(HOGB200_BAD_ARG, "model_read: cellSize missing");
    g.cell_w = v[0]; g.cell_h = v[1];
    if (!get_ints(node, "nbins", v, 1)) return hogk::storage_fail(HOGB200_BAD_ARG, "model_read: nbins missing");
    g.nbins = v[0];
    m->win_sigma = node.count("winSigma") && !node["winSigma"].empty() ? text_real(node["winSigma"][0]) : -1.;
    m->l2hys_threshold = node.count("L2HysThreshold") && !node["L2HysThreshold"].empty() ? text_real(node["L2HysThreshold"][0]) : 0.2;
    m->gamma_correction = get_ints(node, "gammaCorrection", v, 1) ? (v[0] != 0) : 1;
    m->nlevels = get_ints(node, "nlevels", v, 1) ? v[0] : 64;
    // what this implementation (like cv::cuda::HOG) cannot honour is refused rather than ignored
    if (get_ints(node, "histogramNormType", v, 1) && v[0] != 0)
        return hogk::storage_fail(HOGB200_BAD_ARG, "model_read: only histogramNormType L2Hys (0) is supported");
    if (get_ints(node, "signedGradient", v, 1) && v[0] != 0)
        return hogk::storage_fail(HOGB200_BAD_ARG, "model_read: signedGradient is not supported");
    if (get_ints(node, "derivAperture", v, 1) && v[0] != 1 && v[0] != 0)
        return hogk::storage_fail(HOGB200_BAD_ARG, "model_read: only derivAperture 1 is supported");
    Node::const_iterator d = node.find("SVMDetector");
    const size_t n = d == node.end() ? 0 : d->second.size();
    if (count) *count = n;
    if (n > capacity) return hogk::storage_fail(HOGB200_CAPACITY, "model_read: detector capacity too small");
    if (n && !detector) return hogk::storage_fail(HOGB200_BAD_ARG, "model_read: null detector buffer");
    for (size_t i = 0; i < n; i++) detector[i] = (float)text_real(d->second[i]);
    return HOGB200_OK;
}

}  // extern "C"
