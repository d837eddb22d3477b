// ITMSwappingEngine_B200.h -- drop-in subclass of the reference's swapping-engine interface
// (ITMLib/Engine/ITMSwappingEngine.h:21-32) on libdsb200's dsb200_swap_* entry points.  Replaces
// ITMSwappingEngine_CUDA<TVoxel,ITMVoxelBlockHash> in ITMLib/Engine/ITMDenseMapper.cpp:20-33; DenseSLAM reaches it through
// GetSwappingEngine()->SaveToGlobalMemory(scene) when it parks a local map in host memory (DenseSLAM/DenseSlam.h:248-251).
//
// The reference keeps the swap states and the host block store in scene->globalCache and lets AllocateSceneFromDepth update
// the states; here each scene gets a dsb200_swap (device states, pinned host store) owned by this engine, and the
// allocator's swap-state pass (CUDA/ITMSceneReconstructionEngine_CUDA.cu:1730-1733) runs at the start of
// IntegrateGlobalIntoLocal, the first reader of those states after an allocation.
#pragma once

#include <map>

#include "ITMLib/Engine/ITMSwappingEngine.h"
#include "ITMSceneReconstructionEngine_B200.h"

namespace ITMLib
{
	namespace Engine
	{
		template<class TVoxel, class TIndex>
		class ITMSwappingEngine_B200;

		template<class TVoxel>
		class ITMSwappingEngine_B200<TVoxel, ITMVoxelBlockHash> : public ITMSwappingEngine<TVoxel, ITMVoxelBlockHash>
		{
		private:
			void *stream;
			dsb200_recon *recon;                       // told when the volume changed behind it (may be NULL)
			std::map<const void*, dsb200_swap*> caches; // one ITMGlobalCache equivalent per scene

			dsb200_swap *cacheOf(ITMScene<TVoxel, ITMVoxelBlockHash> *scene)
			{
				dsb200_swap *&c = caches[scene];
				if (c == NULL)
				{
					b200::check(dsb200_swap_create(&c, SDF_BUCKET_NUM + SDF_EXCESS_LIST_SIZE, SDF_TRANSFER_BLOCK_NUM, stream));
					b200::check(dsb200_swap_attach_recon(c, recon));   // the library syncs the engine before / notifies it after every swap call
				}
				return c;
			}

			void changed() { }   // done inside libdsb200 since dsb200_swap_attach_recon

		public:
			void IntegrateGlobalIntoLocal(ITMScene<TVoxel, ITMVoxelBlockHash> *scene, ITMRenderState *renderState)
			{
				dsb200_swap *c = cacheOf(scene);
				dsb200_scene s = b200::wrap(scene);
				dsb200_render_state r = b200::wrap(static_cast<ITMRenderState_VH*>(renderState));
				b200::check(dsb200_swap_mark_visible(c, &r));
				b200::check(dsb200_swap_integrate_global_into_local(c, &s, NULL));
				changed();
			}

			void SaveToGlobalMemory(ITMScene<TVoxel, ITMVoxelBlockHash> *scene, ITMRenderState *renderState)
			{
				dsb200_scene s = b200::wrap(scene);
				dsb200_render_state r = b200::wrap(static_cast<ITMRenderState_VH*>(renderState));
				b200::check(dsb200_swap_save_to_global_memory(cacheOf(scene), &s, &r, NULL));
				b200::unwrap(s, scene);
				changed();
			}

			void SaveToGlobalMemory(ITMScene<TVoxel, ITMVoxelBlockHash> *scene)
			{
				dsb200_scene s = b200::wrap(scene);
				b200::check(dsb200_swap_save_to_global_memory(cacheOf(scene), &s, NULL, NULL));
				b200::unwrap(s, scene);
				changed();
			}

			// ITMGlobalCache::SaveToFile / ReadFromFile of a scene's store
			void SaveToFile(ITMScene<TVoxel, ITMVoxelBlockHash> *scene, const char *fileName) { b200::check(dsb200_swap_save_file(cacheOf(scene), fileName)); }
			void ReadFromFile(ITMScene<TVoxel, ITMVoxelBlockHash> *scene, const char *fileName) { b200::check(dsb200_swap_load_file(cacheOf(scene), fileName)); }

			explicit ITMSwappingEngine_B200(dsb200_recon *reconEngine = NULL, void *cudaStream = NULL) : stream(cudaStream), recon(reconEngine) { }
			~ITMSwappingEngine_B200(void)
			{
				for (typename std::map<const void*, dsb200_swap*>::iterator it = caches.begin(); it != caches.end(); ++it) dsb200_swap_destroy(it->second);
			}
		};
	}
}
