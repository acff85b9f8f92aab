#include <cuda_runtime.h>
#include <stdio.h>
__global__ void body(int *c, cudaGraphConditionalHandle h) { int v = atomicAdd(c, 1); cudaGraphSetConditional(h, v < 5 ? 1u : 0u); }
int main() {
	cudaGraph_t g; cudaGraphCreate(&g, 0);
	cudaGraphConditionalHandle h; cudaGraphConditionalHandleCreate(&h, g, 1, cudaGraphCondAssignDefault);
	cudaGraphNodeParams p = { cudaGraphNodeTypeConditional };
	p.conditional.handle = h; p.conditional.type = cudaGraphCondTypeWhile; p.conditional.size = 1;
	cudaGraphNode_t n; cudaGraphAddNode(&n, g, nullptr, 0, &p);
	cudaGraph_t bg = p.conditional.phGraph_out[0];
	int *c; cudaMalloc(&c, 4); cudaMemset(c, 0, 4);
	cudaStream_t s; cudaStreamCreate(&s);
	cudaStreamBeginCaptureToGraph(s, bg, nullptr, nullptr, 0, cudaStreamCaptureModeRelaxed);
	body<<<1,1,0,s>>>(c, h);
	cudaStreamEndCapture(s, nullptr);
	cudaGraphExec_t e; cudaGraphInstantiate(&e, g, 0);
	cudaGraphLaunch(e, s); cudaStreamSynchronize(s);
	int hc; cudaMemcpy(&hc, c, 4, cudaMemcpyDeviceToHost); printf("%d %s\n", hc, cudaGetErrorString(cudaGetLastError()));
}
