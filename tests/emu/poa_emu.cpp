// TEST INFRASTRUCTURE ONLY.  Runtime of the SIMT emulator declared in include/poa_emu.hpp: fibers, warp
// collectives, the kernel launcher and the MUFU.RCP table.
#include "poa_emu.hpp"

#include <map>
#include <mutex>
#include <string>
#include <vector>

uint3 threadIdx, blockIdx;
dim3 blockDim, gridDim;

extern "C" void poa_emu_switch(void** from_rsp, void** to_rsp);
asm(R"(
.text
.globl poa_emu_switch
.type poa_emu_switch,@function
poa_emu_switch:
    pushq %rbp
    pushq %rbx
    pushq %r12
    pushq %r13
    pushq %r14
    pushq %r15
    movq %rsp, (%rdi)
    movq (%rsi), %rsp
    popq %r15
    popq %r14
    popq %r13
    popq %r12
    popq %rbx
    popq %rbp
    ret
.size poa_emu_switch,.-poa_emu_switch
)");

namespace poa_emu
{

namespace
{
constexpr size_t kStackBytes = 256 * 1024;

struct Collective
{
    unsigned mask    = 0;
    unsigned arrived = 0;
    int op           = -1;
    int line         = 0;
    uint64_t gen     = 0;
    uint32_t in[32];
    int arg[32];
    int width[32];
    uint32_t out[32];
};

struct Lane
{
    void* rsp       = nullptr;
    uint8_t* stack  = nullptr;
    bool done       = true;
    int waiting_line = 0;
};
} // namespace

struct Machine
{
    Lane lanes[32];
    void* sched_rsp = nullptr;
    int cur         = -1;
    const std::function<void()>* body = nullptr;
    std::vector<Collective> colls;
    uint8_t* smem   = nullptr;
    size_t smem_sz  = 0;
    uint64_t events = 0; // collectives completed + lanes finished: progress indicator
    unsigned warp_in_block = 0;
    std::vector<float> rcp;
    bool rcp_loaded = false;
};

Machine& machine()
{
    static Machine m;
    return m;
}

int lane_id() { return machine().cur; }
uint8_t* smem_base() { return machine().smem; }
size_t smem_size() { return machine().smem_sz; }

[[noreturn]] void fail(const char* what, int line)
{
    Machine& m = machine();
    std::fprintf(stderr, "poa_emu: %s (line %d), block %u warp %u lane %d\n", what, line, blockIdx.x, m.warp_in_block, m.cur);
    for (int l = 0; l < 32; l++)
        std::fprintf(stderr, "  lane %2d: %s, waiting at line %d\n", l, m.lanes[l].done ? "done" : "live", m.lanes[l].waiting_line);
    std::fflush(stderr);
    std::abort();
}

static void yield_to_scheduler()
{
    Machine& m = machine();
    Lane& me   = m.lanes[m.cur];
    poa_emu_switch(&me.rsp, &m.sched_rsp);
}

static void lane_entry()
{
    Machine& m = machine();
    (*m.body)();
    Lane& me = m.lanes[m.cur];
    me.done  = true;
    m.events++;
    poa_emu_switch(&me.rsp, &m.sched_rsp);
    std::abort(); // a finished lane is never resumed
}

static Collective& coll_for(unsigned mask)
{
    Machine& m = machine();
    for (Collective& c : m.colls)
        if (c.mask == mask)
            return c;
    m.colls.emplace_back();
    m.colls.back().mask = mask;
    return m.colls.back();
}

static void finish(Collective& c)
{
    unsigned ballot = 0;
    if (c.op == kOpBallot)
        for (int l = 0; l < 32; l++)
            if ((c.mask >> l) & 1u)
                ballot |= (c.in[l] & 1u) << l;
    for (int l = 0; l < 32; l++)
    {
        if (!((c.mask >> l) & 1u))
            continue;
        const int w    = c.width[l];
        const int base = l & ~(w - 1);
        int src        = l;
        switch (c.op)
        {
        case kOpSync: c.out[l] = 0; continue;
        case kOpBallot: c.out[l] = ballot; continue;
        case kOpShflIdx: src = base | (c.arg[l] & (w - 1)); break;
        case kOpShflUp: src = (l - c.arg[l] >= base) ? l - c.arg[l] : l; break;
        case kOpShflDown: src = (l + c.arg[l] < base + w) ? l + c.arg[l] : l; break;
        case kOpShflXor: src = ((l ^ c.arg[l]) < base + w) ? (l ^ c.arg[l]) : l; break;
        default: break;
        }
        // reading from a lane outside the mask is undefined on hardware; here it returns that lane's last value
        c.out[l] = c.in[src & 31];
    }
}

uint32_t collective(unsigned mask, Op op, uint32_t value, int arg, int width, int line)
{
    Machine& m     = machine();
    const int lane = m.cur;
    if (lane < 0)
        fail("warp collective outside a kernel", line);
    if (!((mask >> lane) & 1u))
        fail("lane calls a collective whose mask does not name it", line);
    Collective& c = coll_for(mask);
    if (c.arrived == 0)
    {
        c.op   = op;
        c.line = line;
    }
    else if (c.op != static_cast<int>(op) || c.line != line)
    {
        std::fprintf(stderr, "poa_emu: lanes of one mask (0x%08x) wait at different collectives: line %d vs line %d\n", mask,
                     c.line, line);
        fail("divergent collective", line);
    }
    c.in[lane]    = value;
    c.arg[lane]   = arg;
    c.width[lane] = width;
    c.arrived |= 1u << lane;
    if (c.arrived == mask)
    {
        finish(c);
        c.arrived = 0;
        c.gen++;
        m.events++;
    }
    else
    {
        const uint64_t g            = c.gen;
        m.lanes[lane].waiting_line = line;
        while (c.gen == g)
            yield_to_scheduler();
        m.lanes[lane].waiting_line = 0;
    }
    return c.out[lane];
}

static void run_warp(const std::function<void()>& body, unsigned first_thread)
{
    Machine& m = machine();
    m.body     = &body;
    for (Collective& c : m.colls)
        c.arrived = 0;
    for (int l = 0; l < 32; l++)
    {
        Lane& ln = m.lanes[l];
        if (ln.stack == nullptr)
            ln.stack = static_cast<uint8_t*>(std::malloc(kStackBytes));
        uintptr_t top = reinterpret_cast<uintptr_t>(ln.stack + kStackBytes) & ~uintptr_t(15);
        void** sp     = reinterpret_cast<void**>(top - 64);
        for (int k = 0; k < 6; k++)
            sp[k] = nullptr;
        sp[6]   = reinterpret_cast<void*>(&lane_entry);
        sp[7]   = nullptr;
        ln.rsp  = sp;
        ln.done = false;
        ln.waiting_line = 0;
    }
    int idle_rounds = 0;
    for (;;)
    {
        bool any_live       = false;
        const uint64_t seen = m.events;
        for (int l = 0; l < 32; l++)
        {
            Lane& ln = m.lanes[l];
            if (ln.done)
                continue;
            any_live    = true;
            m.cur       = l;
            threadIdx.x = first_thread + static_cast<unsigned>(l);
            poa_emu_switch(&m.sched_rsp, &ln.rsp);
        }
        m.cur = -1;
        if (!any_live)
            break;
        if (m.events == seen)
        {
            if (++idle_rounds > 4)
            {
                m.cur = 0;
                fail("deadlock: no lane can make progress (a collective is missing some of its lanes)", 0);
            }
        }
        else
            idle_rounds = 0;
    }
}

void launch(dim3 grid, dim3 block, size_t smem_bytes, const std::function<void()>& body)
{
    // one kernel at a time, whatever host thread launches it (the window pipeline's workers): the machine is a singleton
    static std::mutex launch_mutex;
    std::lock_guard<std::mutex> lock(launch_mutex);
    Machine& m = machine();
    if (block.x % 32 != 0)
        fail("block size must be a multiple of 32", 0);
    gridDim  = grid;
    blockDim = block;
    std::vector<uint8_t> smem(smem_bytes + 64);
    m.smem    = smem.data();
    m.smem_sz = smem_bytes;
    for (unsigned b = 0; b < grid.x; b++)
    {
        blockIdx.x = b;
        blockIdx.y = blockIdx.z = 0;
        std::memset(smem.data(), 0xA5, smem.size()); // shared memory starts undefined
        for (unsigned w = 0; w < block.x / 32; w++)
        {
            m.warp_in_block = w;
            run_warp(body, w * 32);
        }
    }
    m.smem    = nullptr;
    m.smem_sz = 0;
}

// MUFU.RCP(float(i)) for i < 65536 as measured on a B200: a .npy file of float32 (tests/golden/b200_mufu_rcp_u16.npy),
// path in POA_EMU_RCP_TABLE.
float rcp_approx(float d)
{
    Machine& m = machine();
    if (!m.rcp_loaded)
    {
        m.rcp_loaded     = true;
        const char* path = std::getenv("POA_EMU_RCP_TABLE");
        FILE* f          = path ? std::fopen(path, "rb") : nullptr;
        if (f != nullptr)
        {
            unsigned char head[10];
            if (std::fread(head, 1, 10, f) == 10 && head[0] == 0x93)
            {
                const size_t hlen = head[8] | (static_cast<size_t>(head[9]) << 8);
                std::fseek(f, static_cast<long>(10 + hlen), SEEK_SET);
                m.rcp.resize(65536);
                const size_t got = std::fread(m.rcp.data(), sizeof(float), m.rcp.size(), f);
                m.rcp.resize(got);
            }
            std::fclose(f);
        }
    }
    const int i = static_cast<int>(d);
    if (static_cast<float>(i) == d && i >= 0 && static_cast<size_t>(i) < m.rcp.size())
        return m.rcp[static_cast<size_t>(i)];
    return 1.0f / d;
}

static std::map<std::string, long long>& counters()
{
    static std::map<std::string, long long> c;
    return c;
}
void count(const char* name, long long n)
{
    if (lane_id() <= 0) // once per warp
        counters()[name] += n;
}

} // namespace poa_emu

// "name=value;..." of the coverage counters since the last call
extern "C" int poa_emu_counters(char* buf, int cap)
{
    std::string out;
    for (const auto& kv : poa_emu::counters())
        out += kv.first + "=" + std::to_string(kv.second) + ";";
    poa_emu::counters().clear();
    if (buf != nullptr && cap > 0)
    {
        std::strncpy(buf, out.c_str(), static_cast<size_t>(cap) - 1);
        buf[cap - 1] = 0;
    }
    return static_cast<int>(out.size());
}
