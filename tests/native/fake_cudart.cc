/*
 * fake_cudart.cc -- TEST-ONLY stand-in for the handful of CUDA runtime calls the host side of
 * libnvme_strom makes, so that the GPU transfer engine (staging ring, copy stream, events,
 * completion) can run under ThreadSanitizer / AddressSanitizer on a machine without a GPU.
 * Linked into tests/native/gpu_pipeline_san.cc by tests/test_sanitizers_cpu.py and nowhere
 * else: the shipped library links the real libcudart and has no such path.
 *
 * "Device memory" is host memory handed out by fake_device_alloc(); a stream is a thread that
 * executes its queue in order (copies sleep a little first, so that completion really is
 * asynchronous); an event is complete when the queue has passed its record.
 */
#include <cuda_runtime_api.h>

#include <atomic>
#include <chrono>
#include <condition_variable>
#include <cstdlib>
#include <cstring>
#include <deque>
#include <mutex>
#include <thread>
#include <vector>

namespace {

struct FakeEvent {
	std::mutex m;
	std::condition_variable cv;
	uint64_t recorded = 0, completed = 0;
};

struct Op {
	void *dst = nullptr;
	const void *src = nullptr;
	size_t n = 0;
	FakeEvent *ev = nullptr;
	uint64_t gen = 0;
};

struct FakeStream {
	std::mutex m;
	std::condition_variable cv;
	std::deque<Op> q;
	std::thread thr;
	FakeStream() : thr([this] { run(); }) { thr.detach(); }
	void run()
	{
		for (;;)
		{
			Op op;
			{
				std::unique_lock<std::mutex> g(m);
				cv.wait(g, [&] { return !q.empty(); });
				op = q.front();
				q.pop_front();
			}
			if (op.ev)
			{
				std::lock_guard<std::mutex> g(op.ev->m);
				op.ev->completed = op.gen;
				op.ev->cv.notify_all();
			}
			else
			{
				std::this_thread::sleep_for(std::chrono::microseconds(20 + op.n / 50000));
				memcpy(op.dst, op.src, op.n);
			}
		}
	}
	void push(const Op &op)
	{
		{
			std::lock_guard<std::mutex> g(m);
			q.push_back(op);
		}
		cv.notify_one();
	}
};

std::mutex g_ranges_lock;
std::vector<std::pair<char *, size_t>> g_ranges;
std::atomic<long> g_copies{0}, g_copy_bytes{0};

} /* namespace */

extern "C" void *fake_device_alloc(size_t n)
{
	void *p = nullptr;
	if (posix_memalign(&p, 65536, n) != 0)
		return nullptr;
	memset(p, 0xEE, n);
	std::lock_guard<std::mutex> g(g_ranges_lock);
	g_ranges.emplace_back(static_cast<char *>(p), n);
	return p;
}
extern "C" long fake_copy_count() { return g_copies.load(); }
extern "C" long fake_copy_bytes() { return g_copy_bytes.load(); }

extern "C" {

cudaError_t CUDARTAPI cudaGetLastError(void) { return cudaSuccess; }
const char *CUDARTAPI cudaGetErrorName(cudaError_t) { return "fakeError"; }
const char *CUDARTAPI cudaGetErrorString(cudaError_t) { return "fake error"; }
cudaError_t CUDARTAPI cudaSetDevice(int device) { return device == 0 ? cudaSuccess : cudaErrorInvalidDevice; }

cudaError_t CUDARTAPI cudaGetDriverEntryPoint(const char *, void **funcPtr, unsigned long long,
											  enum cudaDriverEntryPointQueryResult *driverStatus)
{
	if (funcPtr) *funcPtr = nullptr;
	if (driverStatus) *driverStatus = cudaDriverEntryPointSymbolNotFound;
	return cudaErrorNotSupported;            /* no driver API here: the library falls back to cudaSetDevice */
}

cudaError_t CUDARTAPI cudaPointerGetAttributes(struct cudaPointerAttributes *attr, const void *ptr)
{
	memset(attr, 0, sizeof(*attr));
	attr->type = cudaMemoryTypeUnregistered;
	std::lock_guard<std::mutex> g(g_ranges_lock);
	for (auto &r : g_ranges)
		if ((const char *)ptr >= r.first && (const char *)ptr < r.first + r.second)
		{
			attr->type = cudaMemoryTypeDevice;
			attr->device = 0;
			attr->devicePointer = const_cast<void *>(ptr);
		}
	return cudaSuccess;
}

cudaError_t CUDARTAPI cudaStreamCreateWithFlags(cudaStream_t *s, unsigned int)
{
	*s = reinterpret_cast<cudaStream_t>(new FakeStream());
	return cudaSuccess;
}

cudaError_t CUDARTAPI cudaHostAlloc(void **p, size_t n, unsigned int)
{
	return posix_memalign(p, 4096, n) == 0 ? cudaSuccess : cudaErrorMemoryAllocation;
}
cudaError_t CUDARTAPI cudaHostRegister(void *, size_t, unsigned int) { return cudaSuccess; }
cudaError_t CUDARTAPI cudaHostUnregister(void *) { return cudaSuccess; }

cudaError_t CUDARTAPI cudaEventCreateWithFlags(cudaEvent_t *e, unsigned int)
{
	*e = reinterpret_cast<cudaEvent_t>(new FakeEvent());
	return cudaSuccess;
}

cudaError_t CUDARTAPI cudaMemcpyAsync(void *dst, const void *src, size_t n, enum cudaMemcpyKind, cudaStream_t s)
{
	Op op;
	op.dst = dst;
	op.src = src;
	op.n = n;
	g_copies++;
	g_copy_bytes += (long)n;
	reinterpret_cast<FakeStream *>(s)->push(op);
	return cudaSuccess;
}

cudaError_t CUDARTAPI cudaEventRecord(cudaEvent_t e, cudaStream_t s)
{
	FakeEvent *ev = reinterpret_cast<FakeEvent *>(e);
	Op op;
	op.ev = ev;
	{
		std::lock_guard<std::mutex> g(ev->m);
		op.gen = ++ev->recorded;
	}
	reinterpret_cast<FakeStream *>(s)->push(op);
	return cudaSuccess;
}

cudaError_t CUDARTAPI cudaEventSynchronize(cudaEvent_t e)
{
	FakeEvent *ev = reinterpret_cast<FakeEvent *>(e);
	std::unique_lock<std::mutex> g(ev->m);
	const uint64_t want = ev->recorded;
	ev->cv.wait(g, [&] { return ev->completed >= want; });
	return cudaSuccess;
}

cudaError_t CUDARTAPI cudaEventQuery(cudaEvent_t e)
{
	FakeEvent *ev = reinterpret_cast<FakeEvent *>(e);
	std::lock_guard<std::mutex> g(ev->m);
	return ev->completed >= ev->recorded ? cudaSuccess : cudaErrorNotReady;
}

cudaError_t CUDARTAPI cudaEventDestroy(cudaEvent_t e)
{
	delete reinterpret_cast<FakeEvent *>(e);
	return cudaSuccess;
}

/* the fake stream's thread is detached and idles forever: nothing to take down */
cudaError_t CUDARTAPI cudaStreamDestroy(cudaStream_t) { return cudaSuccess; }

cudaError_t CUDARTAPI cudaFreeHost(void *p)
{
	free(p);
	return cudaSuccess;
}

} /* extern "C" */
