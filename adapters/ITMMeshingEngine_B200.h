// ITMMeshingEngine_B200.h -- drop-in subclass of the reference's meshing-engine interface
// (ITMLib/Engine/ITMMeshingEngine.h:19-26) forwarding to libdsb200's dsb200_mesh_scene.  Replaces
// ITMMeshingEngine_CUDA<TVoxel,ITMVoxelBlockHash> in ITMLib/Engine/ITMMainEngine.cpp:57-81 (meshingEngine, used by
// ITMMainEngine::SaveSceneToMesh, ITMMainEngine.cpp:161-176).  The mesh stays the reference's ITMMesh (MEMORYDEVICE_CUDA);
// unlike the reference CUDA engine the triangle order is the CPU engine's (deterministic).
#pragma once

#include "ITMLib/Engine/ITMMeshingEngine.h"
#include "ITMSceneReconstructionEngine_B200.h"

namespace ITMLib
{
	namespace Engine
	{
		template<class TVoxel, class TIndex>
		class ITMMeshingEngine_B200;

		template<class TVoxel>
		class ITMMeshingEngine_B200<TVoxel, ITMVoxelBlockHash> : public ITMMeshingEngine<TVoxel, ITMVoxelBlockHash>
		{
		private:
			dsb200_mesher *engine;

		public:
			explicit ITMMeshingEngine_B200(void *cudaStream = NULL) : engine(NULL)
			{
				b200::check(dsb200_mesher_create(&engine, SDF_LOCAL_BLOCK_NUM, cudaStream));
			}
			~ITMMeshingEngine_B200(void) { dsb200_mesher_destroy(engine); }

			void MeshScene(ITMMesh *mesh, const ITMScene<TVoxel, ITMVoxelBlockHash> *scene)
			{
				static_assert(sizeof(ITMMesh::Triangle) == sizeof(dsb200_triangle), "ITMMesh::Triangle layout");
				dsb200_scene s = b200::wrap(const_cast<ITMScene<TVoxel, ITMVoxelBlockHash>*>(scene));
				uint32_t n = 0;
				b200::check(dsb200_mesh_scene(engine, &s, reinterpret_cast<dsb200_triangle*>(mesh->triangles->GetData(MEMORYDEVICE_CUDA)),
					ITMMesh::noMaxTriangles, &n));
				mesh->noTotalTriangles = n;
			}
		};
	}
}
