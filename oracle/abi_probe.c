/*
 * abi_probe.c -- TEST INFRASTRUCTURE.  Compiled against the REFERENCE's own
 * header (-I/root/reference/kmod, file kmod/nvme_strom.h:17-165) and prints
 * its ioctl numbers, struct sizes and field offsets as JSON.  The output is
 * committed as tests/golden/abi_layout.json (see oracle/Makefile, target
 * "golden") and pins include/nvme_strom.h to the reference ABI.
 */
#include <stdio.h>
#include <stddef.h>
#include "nvme_strom.h"

#define S(T)      printf("  \"sizeof:" #T "\": %zu,\n", sizeof(T))
#define F(T, f)   printf("  \"offsetof:" #T "." #f "\": %zu,\n", offsetof(T, f))
#define C(name)   printf("  \"ioctl:" #name "\": %u,\n", (unsigned)(name))

int main(void)
{
	printf("{\n");
	C(STROM_IOCTL__CHECK_FILE); C(STROM_IOCTL__MAP_GPU_MEMORY);
	C(STROM_IOCTL__UNMAP_GPU_MEMORY); C(STROM_IOCTL__LIST_GPU_MEMORY);
	C(STROM_IOCTL__INFO_GPU_MEMORY); C(STROM_IOCTL__ALLOC_DMA_BUFFER);
	C(STROM_IOCTL__MEMCPY_SSD2GPU); C(STROM_IOCTL__MEMCPY_SSD2RAM);
	C(STROM_IOCTL__MEMCPY_WAIT); C(STROM_IOCTL__STAT_INFO);

	S(StromCmd__CheckFile);
	F(StromCmd__CheckFile, fdesc); F(StromCmd__CheckFile, numa_node_id);
	F(StromCmd__CheckFile, support_dma64);
	S(StromCmd__MapGpuMemory);
	F(StromCmd__MapGpuMemory, handle); F(StromCmd__MapGpuMemory, gpu_page_sz);
	F(StromCmd__MapGpuMemory, gpu_npages); F(StromCmd__MapGpuMemory, vaddress);
	F(StromCmd__MapGpuMemory, length);
	S(StromCmd__UnmapGpuMemory); F(StromCmd__UnmapGpuMemory, handle);
	S(StromCmd__ListGpuMemory);
	F(StromCmd__ListGpuMemory, nrooms); F(StromCmd__ListGpuMemory, nitems);
	F(StromCmd__ListGpuMemory, handles);
	S(StromCmd__InfoGpuMemory);
	F(StromCmd__InfoGpuMemory, handle); F(StromCmd__InfoGpuMemory, nrooms);
	F(StromCmd__InfoGpuMemory, nitems); F(StromCmd__InfoGpuMemory, version);
	F(StromCmd__InfoGpuMemory, gpu_page_sz); F(StromCmd__InfoGpuMemory, owner);
	F(StromCmd__InfoGpuMemory, map_offset); F(StromCmd__InfoGpuMemory, map_length);
	F(StromCmd__InfoGpuMemory, paddrs);
	S(StromCmd__MemCopySsdToGpu);
	F(StromCmd__MemCopySsdToGpu, dma_task_id); F(StromCmd__MemCopySsdToGpu, nr_ram2gpu);
	F(StromCmd__MemCopySsdToGpu, nr_ssd2gpu); F(StromCmd__MemCopySsdToGpu, nr_dma_submit);
	F(StromCmd__MemCopySsdToGpu, nr_dma_blocks); F(StromCmd__MemCopySsdToGpu, handle);
	F(StromCmd__MemCopySsdToGpu, offset); F(StromCmd__MemCopySsdToGpu, file_desc);
	F(StromCmd__MemCopySsdToGpu, nr_chunks); F(StromCmd__MemCopySsdToGpu, chunk_sz);
	F(StromCmd__MemCopySsdToGpu, relseg_sz); F(StromCmd__MemCopySsdToGpu, chunk_ids);
	F(StromCmd__MemCopySsdToGpu, wb_buffer);
	S(StromCmd__MemCopyWait);
	F(StromCmd__MemCopyWait, dma_task_id); F(StromCmd__MemCopyWait, status);
	S(StromCmd__MemCopySsdToRam);
	F(StromCmd__MemCopySsdToRam, dma_task_id); F(StromCmd__MemCopySsdToRam, nr_ram2ram);
	F(StromCmd__MemCopySsdToRam, nr_ssd2ram); F(StromCmd__MemCopySsdToRam, nr_dma_submit);
	F(StromCmd__MemCopySsdToRam, nr_dma_blocks); F(StromCmd__MemCopySsdToRam, dest_uaddr);
	F(StromCmd__MemCopySsdToRam, file_desc); F(StromCmd__MemCopySsdToRam, nr_chunks);
	F(StromCmd__MemCopySsdToRam, chunk_sz); F(StromCmd__MemCopySsdToRam, relseg_sz);
	F(StromCmd__MemCopySsdToRam, chunk_ids);
	S(StromCmd__AllocDMABuffer);
	F(StromCmd__AllocDMABuffer, length); F(StromCmd__AllocDMABuffer, node_id);
	F(StromCmd__AllocDMABuffer, dmabuf_fdesc);
	S(StromCmd__StatInfo);
	F(StromCmd__StatInfo, version); F(StromCmd__StatInfo, has_debug);
	F(StromCmd__StatInfo, tsc); F(StromCmd__StatInfo, nr_ssd2gpu);
	F(StromCmd__StatInfo, clk_ssd2gpu); F(StromCmd__StatInfo, nr_setup_prps);
	F(StromCmd__StatInfo, clk_setup_prps); F(StromCmd__StatInfo, nr_submit_dma);
	F(StromCmd__StatInfo, clk_submit_dma); F(StromCmd__StatInfo, nr_wait_dtask);
	F(StromCmd__StatInfo, clk_wait_dtask); F(StromCmd__StatInfo, nr_wrong_wakeup);
	F(StromCmd__StatInfo, cur_dma_count); F(StromCmd__StatInfo, max_dma_count);
	F(StromCmd__StatInfo, nr_debug1); F(StromCmd__StatInfo, clk_debug1);
	F(StromCmd__StatInfo, nr_debug4); F(StromCmd__StatInfo, clk_debug4);
	printf("  \"pathname\": \"%s\"\n}\n", NVME_STROM_IOCTL_PATHNAME);
	return 0;
}
