PatchCollection = None
