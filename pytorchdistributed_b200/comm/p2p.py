"""placeholder replaced below"""
def shutdown():
    pass
