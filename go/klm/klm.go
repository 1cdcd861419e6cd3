// Package klm binds libklm.so (include/klm.h) for the k8s-llm-monitor Go host.
//
// NOT COMPILED IN THIS REPOSITORY'S BUILD ENVIRONMENT (no Go toolchain there);
// the same C ABI is exercised by the ctypes tests. See INTEGRATION.md for the
// call-site changes in internal/scheduler/controller.go and
// internal/metrics/manager.go that use it.
package klm

/*
#cgo CFLAGS: -I${SRCDIR}/../../include
#cgo LDFLAGS: -lklm -lcudart
#include "klm.h"
*/
import "C"

import (
	"fmt"
	"runtime"
	"unsafe"
)

// Node flags: which of buildCandidates' per-item filters a node passes
// (internal/scheduler/controller.go:190, :198).
const (
	NodeHasName  = 1
	NodeStatusOK = 2
)

// Sharding modes of a handle over several GPUs (KLM_SHARD_*): where the node table lives.
const (
	ShardRequests = 0x000 // every device holds the whole node table
	ShardNodes    = 0x100 // device g holds nodes [g L, (g+1) L); peers read them over NVLink
)

// Handle drives one GPU, or several (NewMulti). All methods are safe for concurrent use.
type Handle struct{ h *C.klm_handle }

// klm_last_error() is thread-local and a goroutine may move to another OS thread between
// two cgo calls, so every wrapper runs its call and the error fetch on one locked thread.
func call(what string, f func() C.int) error {
	runtime.LockOSThread()
	defer runtime.UnlockOSThread()
	if rc := f(); rc != 0 {
		return fmt.Errorf("%s: %s", what, C.GoString(C.klm_last_error()))
	}
	return nil
}

// New opens device `device`. There is no CPU fallback: on a machine without a
// B200 this returns an error and the caller keeps its Go code path.
func New(device int) (*Handle, error) { return NewMulti([]int{device}) }

// NewMulti opens one handle over several GPUs: NodesSet / ScheduleBatch / Reconcile then
// spread every batch over all of them (cmd/scheduler/main.go:55-62 builds one controller
// per process, so one handle per process is the natural shape).
func NewMulti(devices []int) (*Handle, error) {
	if len(devices) == 0 {
		return nil, fmt.Errorf("klm: no device")
	}
	ids := make([]C.int, len(devices))
	for i, d := range devices {
		ids[i] = C.int(d)
	}
	var h *C.klm_handle
	if err := call("klm_create", func() C.int { return C.klm_create(&ids[0], C.int(len(ids)), &h) }); err != nil {
		return nil, err
	}
	return &Handle{h}, nil
}

func (k *Handle) Close() { C.klm_destroy(k.h); k.h = nil }

func f64p(s []float64) *C.double {
	if len(s) == 0 {
		return nil
	}
	return (*C.double)(unsafe.Pointer(&s[0]))
}
func f32p(s []float32) *C.float {
	if len(s) == 0 {
		return nil
	}
	return (*C.float)(unsafe.Pointer(&s[0]))
}
func i64p(s []int64) *C.int64_t {
	if len(s) == 0 {
		return nil
	}
	return (*C.int64_t)(unsafe.Pointer(&s[0]))
}
func i32p(s []int32) *C.int32_t {
	if len(s) == 0 {
		return nil
	}
	return (*C.int32_t)(unsafe.Pointer(&s[0]))
}
func u8p(s []uint8) *C.uint8_t {
	if len(s) == 0 {
		return nil
	}
	return (*C.uint8_t)(unsafe.Pointer(&s[0]))
}

// NodesSet installs the UAV snapshot of one reconcile (controller.go:96-98).
func (k *Handle) NodesSet(battery []float64, flags []uint8) error {
	if len(battery) != len(flags) {
		return fmt.Errorf("klm: battery and flags differ in length")
	}
	if err := call("klm_nodes_set", func() C.int { return C.klm_nodes_set(k.h, C.int64_t(len(battery)), f64p(battery), u8p(flags)) }); err != nil {
		return err
	}
	return nil
}

// ScheduleBatch replaces buildCandidates + sort.SliceStable + candidates[0]
// (controller.go:152-161, :174-221) for a batch of independent requests.
// idx[r] == -1 means "no candidate" (controller.go:153-158).
func (k *Handle) ScheduleBatch(minBattery []float64, prefOff, prefIDs []int32) ([]int32, []float64, error) {
	return k.ScheduleBatchMode(minBattery, prefOff, prefIDs, ShardRequests)
}

// ScheduleBatchMode is ScheduleBatch with a sharding mode (ShardRequests / ShardNodes) for a
// handle over several GPUs; on one GPU the modes are the same thing.
func (k *Handle) ScheduleBatchMode(minBattery []float64, prefOff, prefIDs []int32, mode int) ([]int32, []float64, error) {
	r := len(minBattery)
	idx, score := make([]int32, r), make([]float64, r)
	if r == 0 {
		return idx, score, nil
	}
	if len(prefOff) != r+1 {
		return nil, nil, fmt.Errorf("klm: prefOff must have len(minBattery)+1 entries")
	}
	if n := int(prefOff[r]); n < 0 || n > len(prefIDs) {
		return nil, nil, fmt.Errorf("klm: prefOff[len] = %d but %d preferred ids", n, len(prefIDs))
	}
	ids := prefIDs
	if len(ids) == 0 {
		ids = []int32{0}
	}
	if err := call("klm_schedule_batch", func() C.int { return C.klm_schedule_batch(k.h, C.int64_t(r), f64p(minBattery), i32p(prefOff), i32p(ids),
		i32p(idx), f64p(score), C.int(C.KLM_ALGO_DECOMPOSED|mode)) }); err != nil {
		return nil, nil, err
	}
	return idx, score, nil
}

// Reconcile is NodesSet + ScheduleBatch in one call (klm_reconcile): the whole of
// controller.go:88-110 for one cycle, one lock and -- for a small reconcile -- one upload, one
// download and one synchronisation.
func (k *Handle) Reconcile(battery []float64, flags []uint8, minBattery []float64, prefOff, prefIDs []int32) ([]int32, []float64, error) {
	return k.ReconcileMode(battery, flags, minBattery, prefOff, prefIDs, ShardRequests)
}

// ReconcileMode is Reconcile with a sharding mode for a handle over several GPUs.
func (k *Handle) ReconcileMode(battery []float64, flags []uint8, minBattery []float64, prefOff, prefIDs []int32, mode int) ([]int32, []float64, error) {
	r := len(minBattery)
	if len(battery) != len(flags) {
		return nil, nil, fmt.Errorf("klm: battery and flags differ in length")
	}
	if len(prefOff) != r+1 {
		return nil, nil, fmt.Errorf("klm: prefOff must have len(minBattery)+1 entries")
	}
	if n := int(prefOff[r]); n < 0 || n > len(prefIDs) {
		return nil, nil, fmt.Errorf("klm: prefOff[len] = %d but %d preferred ids", n, len(prefIDs))
	}
	idx, score := make([]int32, r), make([]float64, r)
	ids := prefIDs
	if len(ids) == 0 {
		ids = []int32{0}
	}
	err := call("klm_reconcile", func() C.int {
		return C.klm_reconcile(k.h, C.int64_t(len(battery)), f64p(battery), u8p(flags), C.int64_t(r),
			f64p(minBattery), i32p(prefOff), i32p(ids), i32p(idx), f64p(score), C.int(C.KLM_ALGO_DECOMPOSED|mode))
	})
	if err != nil {
		return nil, nil, err
	}
	return idx, score, nil
}

// NodeCols are the columns calculateClusterMetrics reads (manager.go:499-535).
type NodeCols struct {
	CPUCapacity, CPUUsage, MemoryCapacity, MemoryUsage []int64
	GPUCount                                           []int32
	Healthy                                            []uint8
	GPUUsage                                           []float64
}

// ClusterOut is the numeric part of pkg/metrics.ClusterMetrics.
type ClusterOut struct {
	TotalNodes, HealthyNodes, TotalPods, RunningPods int64
	TotalCPU, UsedCPU, TotalMemory, UsedMemory       int64
	TotalGPUs, AvailableGPUs                         int64
	CPUUsageRate, MemoryUsageRate                    float64
	HealthStatus                                     int32  // 0 healthy, 1 warning, 2 critical
	Issues                                           uint32 // bit0 unhealthy nodes, bit1 high CPU, bit2 high memory
}

// ClusterAggregate replaces Manager.calculateClusterMetrics (manager.go:493-565).
func (k *Handle) ClusterAggregate(c *NodeCols, podRunning []uint8) (*ClusterOut, error) {
	var pin runtime.Pinner
	defer pin.Unpin()
	cc := c.toC(&pin)
	var o C.klm_cluster_out
	err := call("klm_cluster_aggregate", func() C.int {
		return C.klm_cluster_aggregate(k.h, &cc, C.int64_t(len(c.CPUCapacity)), u8p(podRunning),
			C.int64_t(len(podRunning)), &o)
	})
	if err != nil {
		return nil, err
	}
	out := clusterFromC(&o)
	return &out, nil
}

// pinSlice pins the backing array of s for the duration of one call. A C struct that holds
// Go pointers may be passed to C only if those pointers are pinned (cgo rules, Go >= 1.21).
func pinSlice[T any](p *runtime.Pinner, s []T) {
	if len(s) > 0 {
		p.Pin(&s[0])
	}
}

func (c *NodeCols) toC(p *runtime.Pinner) C.klm_node_cols {
	pinSlice(p, c.CPUCapacity)
	pinSlice(p, c.CPUUsage)
	pinSlice(p, c.MemoryCapacity)
	pinSlice(p, c.MemoryUsage)
	pinSlice(p, c.GPUCount)
	pinSlice(p, c.Healthy)
	pinSlice(p, c.GPUUsage)
	return C.klm_node_cols{
		cpu_capacity: i64p(c.CPUCapacity), cpu_usage: i64p(c.CPUUsage),
		memory_capacity: i64p(c.MemoryCapacity), memory_usage: i64p(c.MemoryUsage),
		gpu_count: i32p(c.GPUCount), healthy: u8p(c.Healthy),
		gpu_usage: f64p(c.GPUUsage), n_gpu_usage: C.int64_t(len(c.GPUUsage)),
	}
}

func clusterFromC(o *C.klm_cluster_out) ClusterOut {
	return ClusterOut{
		int64(o.total_nodes), int64(o.healthy_nodes), int64(o.total_pods), int64(o.running_pods),
		int64(o.total_cpu), int64(o.used_cpu), int64(o.total_memory), int64(o.used_memory),
		int64(o.total_gpus), int64(o.available_gpus), float64(o.cpu_usage_rate),
		float64(o.memory_usage_rate), int32(o.health_status), uint32(o.issues)}
}

// FleetScan replaces GetHealthyUAVCount and GetLowBatteryUAVs
// (internal/metrics/sources/uav_metrics.go:197-233); indices come back ascending.
func (k *Handle) FleetScan(statusOK []uint8, remainingPercent []float64, threshold float64) (int64, []int32, error) {
	n := len(statusOK)
	low := make([]int32, n+1)
	var healthy, nLow C.int64_t
	if err := call("klm_fleet_scan", func() C.int { return C.klm_fleet_scan(k.h, C.int64_t(n), u8p(statusOK), f64p(remainingPercent), C.double(threshold),
		&healthy, i32p(low), &nLow) }); err != nil {
		return 0, nil, err
	}
	return int64(healthy), low[:int(nLow)], nil
}

// RingCreate allocates nSeries rows of `window` f32 samples in HBM.
func (k *Handle) RingCreate(nSeries int64, window int32) error {
	if err := call("klm_ring_create", func() C.int { return C.klm_ring_create(k.h, C.int64_t(nSeries), C.int32_t(window)) }); err != nil {
		return err
	}
	return nil
}

// RingIngest appends one sample to each listed series (UpdateUAVReport, manager.go:391-449).
func (k *Handle) RingIngest(seriesIDs []int64, values []float32) error {
	if err := call("klm_ring_ingest", func() C.int { return C.klm_ring_ingest(k.h, C.int64_t(len(seriesIDs)), i64p(seriesIDs), f32p(values)) }); err != nil {
		return err
	}
	return nil
}

// RingIngestTick appends one sample to every series (end of Collect, manager.go:286).
func (k *Handle) RingIngestTick(values []float32) error {
	if err := call("klm_ring_ingest_tick", func() C.int { return C.klm_ring_ingest_tick(k.h, f32p(values)) }); err != nil {
		return err
	}
	return nil
}

// WindowStats are min / max / mean / p95 of each series' window.
type WindowStats struct {
	Min, Max, P95 []float32
	Mean          []float64
}

// RingStats computes window statistics of series [s0, s1).
func (k *Handle) RingStats(s0, s1 int64) (*WindowStats, error) {
	n := int(s1 - s0)
	w := &WindowStats{make([]float32, n), make([]float32, n), make([]float32, n), make([]float64, n)}
	if err := call("klm_ring_stats", func() C.int { return C.klm_ring_stats(k.h, C.int64_t(s0), C.int64_t(s1), f32p(w.Min), f32p(w.Max), f64p(w.Mean), f32p(w.P95)) }); err != nil {
		return nil, err
	}
	return w, nil
}

// HostFloat64s returns a slice of n float64 backed by page-locked C memory
// (klm_host_alloc). Arrays passed to the klm_* calls may be ordinary Go
// slices, but the driver stages those synchronously; page-locked arrays make
// the copies asynchronous, which lets klm_collect overlap its two branches and
// klm_schedule_batch pipeline a large batch over PCIe. Release with HostFree.
func HostFloat64s(n int) ([]float64, error) {
	var p unsafe.Pointer
	if err := call("klm_host_alloc", func() C.int { return C.klm_host_alloc(C.size_t(n*8), &p) }); err != nil {
		return nil, err
	}
	return unsafe.Slice((*float64)(p), n), nil
}

// HostFree releases a slice obtained from a Host* allocator.
func HostFree(p unsafe.Pointer) { C.klm_host_free(p) }

// CollectResult is one Manager.Collect cycle: cluster metrics, per-node rates and the
// window statistics of every series. Min / Max / P95 / Mean are views into one
// page-locked block owned by the Collector (valid until the next Collect).
type CollectResult struct {
	Cluster       ClusterOut
	CPURate       []float64
	MemoryRate    []float64
	Min, Max, P95 []float32
	Mean          []float64
}

// Collector holds the page-locked result block for klm_collect's packed output.
type Collector struct {
	k       *Handle
	nSeries int64
	block   unsafe.Pointer
}

// NewCollector prepares the result block for a ring of nSeries series (RingCreate first).
func (k *Handle) NewCollector(nSeries int64) (*Collector, error) {
	var p unsafe.Pointer
	if err := call("klm_host_alloc", func() C.int { return C.klm_host_alloc(C.size_t(C.klm_collect_packed_bytes(C.int64_t(nSeries))), &p) }); err != nil {
		return nil, err
	}
	return &Collector{k: k, nSeries: nSeries, block: p}, nil
}

// Close releases the result block.
func (c *Collector) Close() { C.klm_host_free(c.block); c.block = nil }

// Collect replaces the tail of Manager.Collect (manager.go:195-319): appends `tick` (one
// sample per series, or nil) to the ring and returns calculateClusterMetrics, the node
// rates and the window statistics from one call and one synchronisation.
func (c *Collector) Collect(tick []float32, cols *NodeCols, podRunning []uint8) (*CollectResult, error) {
	n := len(cols.CPUCapacity)
	res := &CollectResult{CPURate: make([]float64, n), MemoryRate: make([]float64, n)}
	var pin runtime.Pinner
	defer pin.Unpin()
	nc := cols.toC(&pin)
	var cl C.klm_cluster_out
	pinSlice(&pin, tick)
	pinSlice(&pin, podRunning)
	pinSlice(&pin, res.CPURate)
	pinSlice(&pin, res.MemoryRate)
	pin.Pin(&nc)
	pin.Pin(&cl)
	in := C.klm_collect_in{tick_values: f32p(tick), nodes: &nc, n_nodes: C.int64_t(n),
		pod_running: u8p(podRunning), n_pods: C.int64_t(len(podRunning))}
	out := C.klm_collect_out{cluster: &cl, cpu_rate: f64p(res.CPURate), memory_rate: f64p(res.MemoryRate),
		win_packed: c.block}
	if err := call("klm_collect", func() C.int { return C.klm_collect(c.k.h, &in, &out) }); err != nil {
		return nil, err
	}
	res.Cluster = clusterFromC(&cl)
	s := int(c.nSeries)
	c4 := uintptr((s*4 + 255) &^ 255)
	res.Min = unsafe.Slice((*float32)(c.block), s)
	res.Max = unsafe.Slice((*float32)(unsafe.Add(c.block, c4)), s)
	res.P95 = unsafe.Slice((*float32)(unsafe.Add(c.block, 2*c4)), s)
	res.Mean = unsafe.Slice((*float64)(unsafe.Add(c.block, 3*c4)), s)
	return res, nil
}

// Pipeline is the asynchronous collector (klm_collect_begin / klm_collect_end): up to two
// cycles in flight, so the tick upload and the result download of one cycle overlap the window
// statistics of the next. Every array a cycle reads or writes after Begin returns lives in
// page-locked C memory owned by the Pipeline (Go memory may not be written by C after a call
// returns); results are valid until the cycle after next is begun.
type Pipeline struct {
	k       *Handle
	nSeries int64
	nNodes  int
	slot    [2]pipeSlot
	begun   uint64
	ended   uint64
}

type pipeSlot struct {
	block   unsafe.Pointer // packed window statistics
	rates   unsafe.Pointer // cpu_rate | memory_rate, nNodes float64 each
	cluster unsafe.Pointer // klm_cluster_out
}

// NewPipeline prepares the two result sets for a ring of nSeries series and snapshots of
// nNodes nodes (RingCreate first).
func (k *Handle) NewPipeline(nSeries int64, nNodes int) (*Pipeline, error) {
	p := &Pipeline{k: k, nSeries: nSeries, nNodes: nNodes}
	for i := range p.slot {
		sizes := []C.size_t{C.size_t(C.klm_collect_packed_bytes(C.int64_t(nSeries))), C.size_t(16*nNodes + 16),
			C.size_t(unsafe.Sizeof(C.klm_cluster_out{}))}
		ptrs := []*unsafe.Pointer{&p.slot[i].block, &p.slot[i].rates, &p.slot[i].cluster}
		for j := range sizes {
			ptr := ptrs[j]
			if err := call("klm_host_alloc", func() C.int { return C.klm_host_alloc(sizes[j], ptr) }); err != nil {
				p.Close()
				return nil, err
			}
		}
	}
	return p, nil
}

// Close releases the result sets (finish every cycle with End first).
func (p *Pipeline) Close() {
	for i := range p.slot {
		C.klm_host_free(p.slot[i].block)
		C.klm_host_free(p.slot[i].rates)
		C.klm_host_free(p.slot[i].cluster)
		p.slot[i] = pipeSlot{}
	}
}

// Begin enqueues one collect cycle and returns; its inputs are consumed before it returns.
func (p *Pipeline) Begin(tick []float32, cols *NodeCols, podRunning []uint8) error {
	if len(cols.CPUCapacity) != p.nNodes {
		return fmt.Errorf("klm: pipeline was prepared for %d nodes", p.nNodes)
	}
	sl := &p.slot[p.begun&1]
	var pin runtime.Pinner
	defer pin.Unpin()
	nc := cols.toC(&pin)
	pinSlice(&pin, tick)
	pinSlice(&pin, podRunning)
	pin.Pin(&nc)
	in := C.klm_collect_in{tick_values: f32p(tick), nodes: &nc, n_nodes: C.int64_t(p.nNodes),
		pod_running: u8p(podRunning), n_pods: C.int64_t(len(podRunning))}
	out := C.klm_collect_out{cluster: (*C.klm_cluster_out)(sl.cluster), cpu_rate: (*C.double)(sl.rates),
		memory_rate: (*C.double)(unsafe.Add(sl.rates, 8*p.nNodes)), win_packed: sl.block}
	if err := call("klm_collect_begin", func() C.int { return C.klm_collect_begin(p.k.h, &in, &out) }); err != nil {
		return err
	}
	p.begun++
	return nil
}

// End completes the oldest cycle begun and returns its results (views into the Pipeline's
// page-locked memory).
func (p *Pipeline) End() (*CollectResult, error) {
	if err := call("klm_collect_end", func() C.int { return C.klm_collect_end(p.k.h) }); err != nil {
		return nil, err
	}
	sl := &p.slot[p.ended&1]
	p.ended++
	s, n := int(p.nSeries), p.nNodes
	c4 := uintptr((s*4 + 255) &^ 255)
	return &CollectResult{
		Cluster:    clusterFromC((*C.klm_cluster_out)(sl.cluster)),
		CPURate:    unsafe.Slice((*float64)(sl.rates), n),
		MemoryRate: unsafe.Slice((*float64)(unsafe.Add(sl.rates, 8*n)), n),
		Min:        unsafe.Slice((*float32)(sl.block), s),
		Max:        unsafe.Slice((*float32)(unsafe.Add(sl.block, c4)), s),
		P95:        unsafe.Slice((*float32)(unsafe.Add(sl.block, 2*c4)), s),
		Mean:       unsafe.Slice((*float64)(unsafe.Add(sl.block, 3*c4)), s),
	}, nil
}

// RingLoad bulk-loads history: rows [s0, s1) each get `count` samples from data.
func (k *Handle) RingLoad(s0, s1 int64, count int32, data []float32) error {
	if int64(len(data)) != (s1-s0)*int64(count) {
		return fmt.Errorf("klm: RingLoad needs (s1-s0)*count samples")
	}
	return call("klm_ring_load", func() C.int {
		return C.klm_ring_load(k.h, C.int64_t(s0), C.int64_t(s1), C.int32_t(count), f32p(data))
	})
}

// NodeRates holds the per-node arithmetic of buildNodeMetrics (sources/node_metrics.go:116-140)
// and IsUnderPressure (pkg/metrics/types.go:159-162).
type NodeRates struct {
	CPURate, MemoryRate, DiskRate []float64
	DiskUsage                     []int64
	UnderPressure                 []uint8
}

// NodeRates replaces the rate formulas of buildNodeMetrics; diskAllocatable may be nil.
func (k *Handle) NodeRates(cpuCap, cpuUse, memCap, memUse, diskCap, diskAlloc []int64) (*NodeRates, error) {
	n := len(cpuCap)
	if len(cpuUse) != n || len(memCap) != n || len(memUse) != n || (diskCap != nil && len(diskCap) != n) ||
		(diskAlloc != nil && len(diskAlloc) != n) {
		return nil, fmt.Errorf("klm: node columns differ in length")
	}
	r := &NodeRates{make([]float64, n), make([]float64, n), make([]float64, n), make([]int64, n), make([]uint8, n)}
	err := call("klm_node_rates", func() C.int {
		return C.klm_node_rates(k.h, C.int64_t(n), i64p(cpuCap), i64p(cpuUse), i64p(memCap), i64p(memUse),
			i64p(diskCap), i64p(diskAlloc), f64p(r.CPURate), f64p(r.MemoryRate), i64p(r.DiskUsage),
			f64p(r.DiskRate), u8p(r.UnderPressure))
	})
	if err != nil {
		return nil, err
	}
	return r, nil
}

// ContainerCols are the per-container columns of buildPodMetrics (sources/pod_metrics.go:105-127);
// the containers of pod p are rows Off[p]..Off[p+1].
type ContainerCols struct {
	Off                                                                    []int32
	CPUUsage, MemoryUsage, CPURequest, CPULimit, MemoryRequest, MemoryLimit []int64
}

// PodRates holds pod totals [6][P] (cpu usage, memory usage, cpu request, cpu limit, memory
// request, memory limit), rates vs limit, utilisation vs request and IsOverLimit.
type PodRates struct {
	Totals                                   []int64
	CPURate, MemoryRate, CPUUtil, MemoryUtil []float64
	OverLimit                                []uint8
}

// PodRates replaces the arithmetic of buildPodMetrics (sources/pod_metrics.go:104-171) and
// GetResourceUtilization / IsOverLimit (pkg/metrics/types.go:165-184).
func (k *Handle) PodRates(c *ContainerCols) (*PodRates, error) {
	p := len(c.Off) - 1
	if p < 0 {
		return nil, fmt.Errorf("klm: Off needs nPods+1 entries")
	}
	nc := int(c.Off[p])
	for _, col := range [][]int64{c.CPUUsage, c.MemoryUsage, c.CPURequest, c.CPULimit, c.MemoryRequest, c.MemoryLimit} {
		if len(col) < nc {
			return nil, fmt.Errorf("klm: container column shorter than Off[nPods]")
		}
	}
	r := &PodRates{make([]int64, 6*p), make([]float64, p), make([]float64, p), make([]float64, p),
		make([]float64, p), make([]uint8, p)}
	err := call("klm_pod_rates", func() C.int {
		return C.klm_pod_rates(k.h, C.int64_t(p), i32p(c.Off), i64p(c.CPUUsage), i64p(c.MemoryUsage),
			i64p(c.CPURequest), i64p(c.CPULimit), i64p(c.MemoryRequest), i64p(c.MemoryLimit), i64p(r.Totals),
			f64p(r.CPURate), f64p(r.MemoryRate), f64p(r.CPUUtil), f64p(r.MemoryUtil), u8p(r.OverLimit))
	})
	if err != nil {
		return nil, err
	}
	return r, nil
}

// NodeAvailable replaces NodeMetrics.GetAvailableResources (pkg/metrics/types.go:151-156).
func (k *Handle) NodeAvailable(cpuCap, cpuUse, memCap, memUse, diskCap, diskUse []int64) (cores, memGB, diskGB []float64, err error) {
	n := len(cpuCap)
	if len(cpuUse) != n || len(memCap) != n || len(memUse) != n || len(diskCap) != n || len(diskUse) != n {
		return nil, nil, nil, fmt.Errorf("klm: node columns differ in length")
	}
	cores, memGB, diskGB = make([]float64, n), make([]float64, n), make([]float64, n)
	err = call("klm_node_available", func() C.int {
		return C.klm_node_available(k.h, C.int64_t(n), i64p(cpuCap), i64p(cpuUse), i64p(memCap), i64p(memUse),
			i64p(diskCap), i64p(diskUse), f64p(cores), f64p(memGB), f64p(diskGB))
	})
	return
}

// NetworkQuality replaces NetworkMetrics.GetQuality (pkg/metrics/types.go:187-199):
// 0 disconnected, 1 excellent, 2 good, 3 fair, 4 poor.
func (k *Handle) NetworkQuality(connected []uint8, rttMs []float64) ([]uint8, error) {
	if len(connected) != len(rttMs) {
		return nil, fmt.Errorf("klm: connected and rtt differ in length")
	}
	q := make([]uint8, len(connected))
	err := call("klm_network_quality", func() C.int {
		return C.klm_network_quality(k.h, C.int64_t(len(q)), u8p(connected), f64p(rttMs), u8p(q))
	})
	return q, err
}

// PodRestarts replaces the RestartCount sum of buildPodMetrics (sources/pod_metrics.go:174-177).
func (k *Handle) PodRestarts(statusOff, restartCount []int32) ([]int32, error) {
	p := len(statusOff) - 1
	if p < 0 || int(statusOff[p]) > len(restartCount) {
		return nil, fmt.Errorf("klm: statusOff / restartCount lengths disagree")
	}
	out := make([]int32, p)
	cnt := restartCount
	if len(cnt) == 0 {
		cnt = []int32{0}
	}
	err := call("klm_pod_restarts", func() C.int {
		return C.klm_pod_restarts(k.h, C.int64_t(p), i32p(statusOff), i32p(cnt), i32p(out))
	})
	return out, err
}

// NodeHealth replaces the condition scan of buildNodeMetrics (sources/node_metrics.go:142-164):
// condType uses KLM_COND_* (0 Ready, 1..4 pressure / network unavailable, 5 other), condStatus
// 0 False, 1 True, 2 Unknown. failing[i] has bit t set for each condition type that made node i
// unhealthy.
func (k *Handle) NodeHealth(condOff []int32, condType, condStatus []uint8) (healthy, failing []uint8, err error) {
	n := len(condOff) - 1
	if n < 0 || int(condOff[n]) > len(condType) || len(condType) != len(condStatus) {
		return nil, nil, fmt.Errorf("klm: condition columns disagree")
	}
	healthy, failing = make([]uint8, n), make([]uint8, n)
	ct, cs := condType, condStatus
	if len(ct) == 0 {
		ct, cs = []uint8{0}, []uint8{0}
	}
	err = call("klm_node_health", func() C.int {
		return C.klm_node_health(k.h, C.int64_t(n), i32p(condOff), u8p(ct), u8p(cs), u8p(healthy), u8p(failing))
	})
	return
}

// ABIVersion is major*1000 + minor of the library's header.
func ABIVersion() int { return int(C.klm_abi_version()) }

// ShardRange is the node range [lo, hi) that device `part` of `parts` owns in ShardNodes mode.
func ShardRange(nNodes int64, parts, part int) (lo, hi int64) {
	var a, b C.int64_t
	C.klm_shard_range(C.int64_t(nNodes), C.int(parts), C.int(part), &a, &b)
	return int64(a), int64(b)
}

// TopKMax is the largest k ScheduleBatchTopK accepts.
func TopKMax() int { return int(C.klm_topk_max()) }

// ScheduleBatchTopK returns, for every request, its k best candidates in the order
// sort.SliceStable leaves them (controller.go:160); the reference keeps only [0] (:161).
// idx and score are [len(minBattery)][k] flattened, padded with -1 / 0.
func (k *Handle) ScheduleBatchTopK(minBattery []float64, prefOff, prefIDs []int32, topK int) ([]int32, []float64, error) {
	r := len(minBattery)
	if len(prefOff) != r+1 {
		return nil, nil, fmt.Errorf("klm: prefOff must have len(minBattery)+1 entries")
	}
	if n := int(prefOff[r]); n < 0 || n > len(prefIDs) {
		return nil, nil, fmt.Errorf("klm: prefOff[len] = %d but %d preferred ids", n, len(prefIDs))
	}
	if topK < 1 || topK > TopKMax() {
		return nil, nil, fmt.Errorf("klm: k must be in [1, %d]", TopKMax())
	}
	idx, score := make([]int32, r*topK), make([]float64, r*topK)
	if r == 0 {
		return idx, score, nil
	}
	ids := prefIDs
	if len(ids) == 0 {
		ids = []int32{0}
	}
	err := call("klm_schedule_batch_topk", func() C.int {
		return C.klm_schedule_batch_topk(k.h, C.int64_t(r), f64p(minBattery), i32p(prefOff), i32p(ids),
			C.int(topK), i32p(idx), f64p(score))
	})
	if err != nil {
		return nil, nil, err
	}
	return idx, score, nil
}
