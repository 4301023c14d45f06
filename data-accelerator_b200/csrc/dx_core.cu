// dx_core.cu — context, device memory pool, string dictionary host operations.
#include "dx_decode.cuh"

// ------------------------------- pool ---------------------------------------------------
static size_t round_size(size_t n) {
    // 512-byte granules up to 1 MiB, 64 KiB granules above (keeps splits of large blocks from littering the free list)
    if (n < 512) return 512;
    const size_t g = n >= (1u << 20) ? 65536 : 512;
    return (n + g - 1) / g * g;
}
void DxPool::add_free(char* p, size_t n) {
    // merge with the free neighbours that lie in the same slab
    auto sl = slabs.upper_bound(p); --sl;
    char* lo = sl->first; char* hi = sl->first + sl->second;
    auto nx = free_by_addr.lower_bound(p);
    if (nx != free_by_addr.end() && nx->first == p + n && nx->first < hi) { n += nx->second; take_free(nx); }
    auto pv = free_by_addr.lower_bound(p);
    if (pv != free_by_addr.begin()) {
        --pv;
        if (pv->first + pv->second == p && pv->first >= lo) { p = pv->first; n += pv->second; take_free(pv); }
    }
    free_by_addr[p] = n;
    free_by_size.emplace(n, p);
}
void DxPool::take_free(std::map<char*, size_t>::iterator it) {
    auto r = free_by_size.equal_range(it->second);
    for (auto k = r.first; k != r.second; ++k) if (k->second == it->first) { free_by_size.erase(k); break; }
    free_by_addr.erase(it);
}
char* DxPool::new_slab(size_t bytes) {
    void* p = nullptr;
    if (cudaMalloc(&p, bytes) != cudaSuccess) { cudaGetLastError(); return nullptr; }
    slabs[(char*)p] = bytes; bytes_reserved += bytes;
    return (char*)p;
}
void* DxPool::alloc(size_t bytes, bool dedicated) {
    const size_t sz = round_size(bytes);
    std::lock_guard<std::mutex> g(mu);
    if (dedicated) {
        // a free block that is a whole slab of a fitting size serves as well as a new allocation
        for (auto it = free_by_size.lower_bound(sz); it != free_by_size.end() && it->first <= sz + sz / 4; ++it) {
            auto sl = slabs.find(it->second);
            if (sl != slabs.end() && sl->second == it->first) { char* p = it->second; const size_t n = it->first; take_free(free_by_addr.find(p)); live[p] = n; return p; }
        }
        char* p = new_slab(sz);
        if (!p) { mu.unlock(); trim(); mu.lock(); p = new_slab(sz); }
        if (!p) return nullptr;
        live[p] = sz;
        return p;
    }
    auto it = free_by_size.lower_bound(sz);
    char* p; size_t have;
    if (it != free_by_size.end()) { p = it->second; have = it->first; take_free(free_by_addr.find(p)); }
    else {
        size_t grow = bytes_reserved / 2;
        if (grow < (64u << 20)) grow = 64u << 20;
        if (grow > ((size_t)4 << 30)) grow = (size_t)4 << 30;
        size_t want = sz > grow ? sz : grow;
        want = (want + (2u << 20) - 1) / (2u << 20) * (2u << 20);
        p = new_slab(want); have = want;
        if (!p) { mu.unlock(); trim(); mu.lock(); p = new_slab(sz); have = sz; }     // memory is tight: exactly what is needed
        if (!p) return nullptr;
    }
    if (have - sz >= 65536 || (have - sz >= 512 && sz < 65536)) { add_free(p + sz, have - sz); have = sz; }
    live[p] = have;
    return p;
}
void DxPool::release(void* p) {
    if (!p) return;
    std::lock_guard<std::mutex> g(mu);
    auto it = live.find(p);
    if (it == live.end()) return;
    const size_t n = it->second;
    live.erase(it);
    add_free((char*)p, n);
}
void DxPool::trim() {
    std::lock_guard<std::mutex> g(mu);
    for (auto sl = slabs.begin(); sl != slabs.end();) {
        auto f = free_by_addr.find(sl->first);
        if (f != free_by_addr.end() && f->second == sl->second) {
            take_free(f); cudaFree(sl->first); bytes_reserved -= sl->second; sl = slabs.erase(sl);
        } else ++sl;
    }
}
DxPool::~DxPool() {
    for (auto& kv : slabs) cudaFree(kv.first);
}

// ------------------------------- dictionary ---------------------------------------------
__global__ void dict_insert_one(DictDev d, const uint8_t* bytes, uint32_t len, int32_t* code_out, unsigned long long* err) {
    uint64_t w0, w1; dict_pack(bytes, len, w0, w1);
    *code_out = dict_get_or_insert(d, w0, w1, len, bytes, err);
}

int DxDict::init(dx_ctx* c, uint32_t capacity_pow2, uint32_t heap_bytes) {
    ctx = c;
    uint32_t cap = 16; while (cap < capacity_pow2) cap <<= 1;
    slots = DevBuf(c, (size_t)cap * sizeof(DictSlot));
    heap = DevBuf(c, heap_bytes);
    counters = DevBuf(c, 64);
    uint32_t max_codes = cap / 2;
    code_off = DevBuf(c, (size_t)max_codes * 4);
    code_len = DevBuf(c, (size_t)max_codes * 4);
    if (!slots.p || !heap.p || !counters.p || !code_off.p || !code_len.p) return c->set_error(DX_ERR_CUDA, "dictionary allocation failed");
    DX_CUDA(c, cudaMemsetAsync(slots.p, 0, (size_t)cap * sizeof(DictSlot), c->stream));
    DX_CUDA(c, cudaMemsetAsync(counters.p, 0, 64, c->stream));
    dev.slots = slots.as<DictSlot>(); dev.cap_mask = cap - 1; dev.max_codes = max_codes;
    dev.heap = heap.as<uint8_t>(); dev.heap_cap = heap_bytes;
    dev.counters = counters.as<uint32_t>();
    dev.code_off = code_off.as<uint32_t>(); dev.code_len = code_len.as<uint32_t>();
    return DX_OK;
}

int DxDict::insert_host(const std::string& s, int32_t* code_out) {
    dx_ctx* c = ctx;
    auto hit = seeded.find(s);
    if (hit != seeded.end()) { if (code_out) *code_out = hit->second; return DX_OK; }
    DevBuf tmp(c, s.size() + 32);
    uint8_t* d = tmp.as<uint8_t>();
    // layout: [0,8) err flags, [8,12) code, [16, ...) bytes
    DX_CUDA(c, cudaMemsetAsync(d, 0, 16, c->stream));
    if (!s.empty()) DX_CUDA(c, cudaMemcpyAsync(d + 16, s.data(), s.size(), cudaMemcpyHostToDevice, c->stream));
    dict_insert_one<<<1, 1, 0, c->stream>>>(dev, d + 16, (uint32_t)s.size(), reinterpret_cast<int32_t*>(d + 8),
                                            reinterpret_cast<unsigned long long*>(d));
    unsigned long long err = 0; int32_t code = -1;
    DX_CUDA(c, cudaMemcpyAsync(&err, d, 8, cudaMemcpyDeviceToHost, c->stream));
    DX_CUDA(c, cudaMemcpyAsync(&code, d + 8, 4, cudaMemcpyDeviceToHost, c->stream));
    DX_CUDA(c, cudaStreamSynchronize(c->stream));
    if (err || code < 0) return c->set_error(DX_ERR_CAPACITY, "string dictionary overflow");
    seeded[s] = code;
    if (code_out) *code_out = code;
    return DX_OK;
}

int DxDict::sync_to_host() {
    dx_ctx* c = ctx;
    uint32_t cnt[4] = {0, 0, 0, 0};
    DX_CUDA(c, cudaMemcpyAsync(cnt, counters.p, 16, cudaMemcpyDeviceToHost, c->stream));
    DX_CUDA(c, cudaStreamSynchronize(c->stream));
    uint32_t n = cnt[0] < dev.max_codes ? cnt[0] : dev.max_codes;
    size_t have = host_strings.size();
    if (n <= have) return DX_OK;
    std::vector<uint32_t> off(n - have), len(n - have);
    DX_CUDA(c, cudaMemcpyAsync(off.data(), dev.code_off + have, (n - have) * 4, cudaMemcpyDeviceToHost, c->stream));
    DX_CUDA(c, cudaMemcpyAsync(len.data(), dev.code_len + have, (n - have) * 4, cudaMemcpyDeviceToHost, c->stream));
    uint32_t heap_used = cnt[1] < dev.heap_cap ? cnt[1] : dev.heap_cap;
    std::vector<uint8_t> hb(heap_used + 1);
    if (heap_used) DX_CUDA(c, cudaMemcpyAsync(hb.data(), dev.heap, heap_used, cudaMemcpyDeviceToHost, c->stream));
    DX_CUDA(c, cudaStreamSynchronize(c->stream));
    for (size_t i = 0; i < n - have; i++) {
        if ((size_t)off[i] + len[i] > heap_used) return c->set_error(DX_ERR_STATE, "dictionary heap out of sync");
        host_strings.emplace_back(reinterpret_cast<const char*>(hb.data()) + off[i], len[i]);
    }
    return DX_OK;
}

__global__ void dict_bulk_insert(DictDev d, const uint8_t* bytes, const uint32_t* offs, uint32_t n, unsigned long long* err) {
    // one thread, in code order: string k gets code k again
    for (uint32_t k = 0; k < n; k++) {
        const uint8_t* b = bytes + offs[k]; const uint32_t len = offs[k + 1] - offs[k];
        uint64_t w0, w1; dict_pack(b, len, w0, w1);
        if (dict_get_or_insert(d, w0, w1, len, b, err) != (int32_t)k) atomicOr(err, 1ull << 40);
    }
}

// many host strings at once, one thread per string (the table's insert is the one K1's lanes use concurrently)
__global__ void dict_insert_many(DictDev d, const uint8_t* bytes, const uint32_t* offs, uint32_t n, int32_t* codes, unsigned long long* err) {
    for (uint32_t k = blockIdx.x * blockDim.x + threadIdx.x; k < n; k += gridDim.x * blockDim.x) {
        const uint8_t* b = bytes + offs[k]; const uint32_t len = offs[k + 1] - offs[k];
        uint64_t w0, w1; dict_pack(b, len, w0, w1);
        codes[k] = dict_get_or_insert(d, w0, w1, len, b, err);
    }
}
int DxDict::insert_host_many(const std::vector<std::string>& strs, std::vector<int32_t>* codes_out) {
    dx_ctx* c = ctx;
    codes_out->assign(strs.size(), -1);
    std::vector<size_t> todo;
    for (size_t k = 0; k < strs.size(); k++) { auto hit = seeded.find(strs[k]); if (hit != seeded.end()) (*codes_out)[k] = hit->second; else todo.push_back(k); }
    if (todo.empty()) return DX_OK;
    for (int attempt = 0; attempt < 8; attempt++) {
        std::vector<uint32_t> offs(todo.size() + 1, 0); std::string blob;
        for (size_t i = 0; i < todo.size(); i++) { blob += strs[todo[i]]; offs[i + 1] = (uint32_t)blob.size(); }
        DevBuf dbytes(c, blob.size() + 16), doffs(c, offs.size() * 4), dcodes(c, todo.size() * 4 + 16), derr(c, 16);
        if (!dbytes.p || !doffs.p || !dcodes.p || !derr.p) return c->set_error(DX_ERR_CUDA, "dictionary insert: allocation failed");
        if (!blob.empty()) DX_CUDA(c, cudaMemcpyAsync(dbytes.p, blob.data(), blob.size(), cudaMemcpyHostToDevice, c->stream));
        DX_CUDA(c, cudaMemcpyAsync(doffs.p, offs.data(), offs.size() * 4, cudaMemcpyHostToDevice, c->stream));
        DX_CUDA(c, cudaMemsetAsync(derr.p, 0, 16, c->stream));
        const unsigned blocks = (unsigned)std::min<size_t>((todo.size() + 127) / 128, 1024);
        dict_insert_many<<<blocks, 128, 0, c->stream>>>(dev, dbytes.as<uint8_t>(), doffs.as<uint32_t>(), (uint32_t)todo.size(), dcodes.as<int32_t>(), derr.as<unsigned long long>());
        std::vector<int32_t> got(todo.size());
        unsigned long long e = 0;
        DX_CUDA(c, cudaMemcpyAsync(got.data(), dcodes.p, todo.size() * 4, cudaMemcpyDeviceToHost, c->stream));
        DX_CUDA(c, cudaMemcpyAsync(&e, derr.p, 8, cudaMemcpyDeviceToHost, c->stream));
        DX_CUDA(c, cudaStreamSynchronize(c->stream));
        bool full = e != 0;
        for (int32_t g : got) if (g < 0) full = true;
        if (!full) {
            for (size_t i = 0; i < todo.size(); i++) { (*codes_out)[todo[i]] = got[i]; seeded[strs[todo[i]]] = got[i]; }
            return DX_OK;
        }
        int rc = grow(); if (rc) return rc;      // every string that fitted keeps its code; the rest is inserted again
    }
    return c->set_error(DX_ERR_CAPACITY, "string dictionary overflow");
}

int DxDict::overflowed(bool* yes) {
    dx_ctx* c = ctx;
    uint32_t cnt[4] = {0, 0, 0, 0};
    DX_CUDA(c, cudaMemcpyAsync(cnt, counters.p, 16, cudaMemcpyDeviceToHost, c->stream));
    DX_CUDA(c, cudaStreamSynchronize(c->stream));
    *yes = cnt[0] > dev.max_codes || cnt[1] > dev.heap_cap;
    return DX_OK;
}

int DxDict::grow() {
    dx_ctx* c = ctx;
    int rc = sync_to_host(); if (rc) return rc;          // every entry that fitted (codes < max_codes, bytes inside the heap)
    const unsigned long long new_cap = ((unsigned long long)dev.cap_mask + 1) * 4, new_heap = (unsigned long long)dev.heap_cap * 4;
    if (new_cap > (1ull << 28) || new_heap > (1ull << 31)) return c->set_error(DX_ERR_CAPACITY, "string dictionary cannot grow further (64 M entries / 2 GiB of text)");
    std::vector<std::string> keep = host_strings;
    std::map<std::string, int32_t> keep_seeded = seeded;
    rc = init(c, (uint32_t)new_cap, (uint32_t)new_heap); if (rc) return rc;
    host_strings = keep; seeded = keep_seeded;
    if (keep.empty()) return DX_OK;
    std::vector<uint32_t> offs(keep.size() + 1, 0); std::string blob;
    for (size_t k = 0; k < keep.size(); k++) { blob += keep[k]; offs[k + 1] = (uint32_t)blob.size(); }
    DevBuf dbytes(c, blob.size() + 16), doffs(c, offs.size() * 4), derr(c, 16);
    if (!dbytes.p || !doffs.p || !derr.p) return c->set_error(DX_ERR_CUDA, "dictionary growth: allocation failed");
    if (!blob.empty()) DX_CUDA(c, cudaMemcpyAsync(dbytes.p, blob.data(), blob.size(), cudaMemcpyHostToDevice, c->stream));
    DX_CUDA(c, cudaMemcpyAsync(doffs.p, offs.data(), offs.size() * 4, cudaMemcpyHostToDevice, c->stream));
    DX_CUDA(c, cudaMemsetAsync(derr.p, 0, 16, c->stream));
    dict_bulk_insert<<<1, 1, 0, c->stream>>>(dev, dbytes.as<uint8_t>(), doffs.as<uint32_t>(), (uint32_t)keep.size(), derr.as<unsigned long long>());
    unsigned long long e = 0;
    DX_CUDA(c, cudaMemcpyAsync(&e, derr.p, 8, cudaMemcpyDeviceToHost, c->stream));
    DX_CUDA(c, cudaStreamSynchronize(c->stream));
    if (e) return c->set_error(DX_ERR_STATE, "dictionary growth: codes changed while re-inserting");
    c->counters["dict_grows"]++;
    return DX_OK;
}

int DxDict::lookup_host(const std::string& s) const {
    for (size_t i = 0; i < host_strings.size(); i++) if (host_strings[i] == s) return (int)i;
    return -1;
}
